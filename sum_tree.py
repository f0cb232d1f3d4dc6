"""Drop-in for the reference's `sum_tree` extension module.

The reference imports its pybind module by its top-level name
(`from sum_tree import SumTreef as SumTree`, hanabi_agents/rlax_dqn/priority_buffer.py:10;
module definition sum_tree/sum_tree/src/sum_tree_py.cc:7-27).  With this repository's root on
`sys.path` that import resolves here and yields the B200 implementation
(`hanabi_agents_b200.sum_tree.SumTreef`: same constructor, the same nine methods, the same
`__repr__`, IndexError on out-of-range reads).  There is no CPU fallback: without the CUDA
library the import of the class raises.
"""
from hanabi_agents_b200.sum_tree import SumTreef  # noqa: F401

__all__ = ["SumTreef"]
