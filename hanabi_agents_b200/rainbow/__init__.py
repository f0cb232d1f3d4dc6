"""Dopamine-style verbs and TF-Rainbow semantics of hanabi_agents/rainbow/ on the same kernels."""
from .rainbow_agent import RainbowAgent, linearly_decaying_epsilon  # noqa: F401
