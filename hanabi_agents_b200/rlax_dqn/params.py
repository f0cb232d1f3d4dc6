"""RlaxRainbowParams with the reference's field names and defaults
(hanabi_agents/rlax_dqn/params.py:5-20).  gin is optional: when it is installed the class is
registered as configurable exactly like the reference's."""
from typing import Callable, List, NamedTuple, Union


class RlaxRainbowParams(NamedTuple):
    """Parameterization class for the Rainbow DQN agent."""

    train_batch_size: int = 256
    target_update_period: int = 500
    discount: float = 0.99
    epsilon: Union[Callable[[int], float], float] = lambda x: 0.1
    learning_rate: float = 0.001
    layers: List[int] = [512]
    use_double_q: bool = True
    use_priority: bool = True
    experience_buffer_size: int = 2**19
    seed: int = 1234
    n_atoms: int = 51
    atom_vmax: int = 25
    beta_is: Union[Callable[[int], float], float] = lambda x: 0.4


try:  # pragma: no cover - gin is not in this image
    import gin
    RlaxRainbowParams = gin.configurable(RlaxRainbowParams)
except ImportError:
    pass
