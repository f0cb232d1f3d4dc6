"""Transition tuple (hanabi_agents/rlax_dqn/transition.py:12-14)."""
from collections import namedtuple

Transition = namedtuple(
    "Transition",
    ["observation_tm1", "action_tm1", "reward_t", "observation_t", "legal_moves_t", "terminal_t"])
