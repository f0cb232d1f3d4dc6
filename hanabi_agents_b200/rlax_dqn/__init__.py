"""Drop-in for hanabi_agents.rlax_dqn (reference: hanabi_agents/rlax_dqn/__init__.py:1-2)."""
from .params import RlaxRainbowParams  # noqa: F401


def __getattr__(name):
    # the agent / buffers need the CUDA library; import them lazily so that the parameter class
    # stays importable on a machine without it
    if name == "DQNAgent":
        from .agent import DQNAgent
        return DQNAgent
    if name in ("ExperienceBuffer", "PriorityBuffer"):
        from . import buffers
        return getattr(buffers, name)
    if name == "Transition":
        from .transition import Transition
        return Transition
    raise AttributeError(name)
