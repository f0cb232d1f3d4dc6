"""Rule-based agents (hanabi_agents/rule_based/__init__.py of the reference)."""
from .rule_based import RulebasedAgent
from .ruleset import Rule, Ruleset
from . import predefined_rules
