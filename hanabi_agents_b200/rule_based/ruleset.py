"""Rule descriptors with the reference's names (hanabi_agents/rule_based/ruleset.py:202-655).

In the reference every rule is a Python function over one HLE observation object; here a rule
is a small descriptor (id + arguments) and the whole rule list is evaluated for every game of
the batch by one CUDA kernel (csrc/hanabi_rules.cuh).  `Ruleset.<name>` keeps the reference's
spelling, the factories keep their signatures.
"""


class Rule:
    __slots__ = ("rule_id", "name", "farg", "iarg")

    def __init__(self, rule_id, name, farg=0.0, iarg=0):
        self.rule_id, self.name, self.farg, self.iarg = int(rule_id), name, float(farg), int(iarg)

    def __repr__(self):
        return "<Rule %s(%g, %d)>" % (self.name, self.farg, self.iarg)

    def __call__(self, observation):
        raise TypeError(
            "rules run batched on the GPU: pass them to RulebasedAgent, which acts on a "
            "HanabiParallelEnv; they are not callable on single observations")


class Ruleset:
    discard_oldest_first = Rule(0, "discard_oldest_first")          # ruleset.py:206
    osawa_discard = Rule(1, "osawa_discard")                        # :220
    tell_unknown = Rule(2, "tell_unknown")                          # :285
    tell_randomly = Rule(3, "tell_randomly")                        # :314
    play_safe_card = Rule(4, "play_safe_card")                      # :350
    play_if_certain = Rule(5, "play_if_certain")                    # :383
    tell_playable_card_outer = Rule(6, "tell_playable_card_outer")  # :413
    tell_anyone_useful_card = Rule(8, "tell_anyone_useful_card")    # :518
    hail_mary = Rule(10, "hail_mary")                               # :653
    legal_random = Rule(11, "legal_random")                         # :598 (intended semantics)
    discard_randomly = Rule(12, "discard_randomly")                 # :607 (intended semantics)
    tell_most_information = Rule(13, "tell_most_information")       # :539 (always None, :562)

    @staticmethod
    def osawa_discard_factory(unknown_color="quirk"):
        """osawa_discard with the reading of ruleset.py:229-230 made explicit.  "quirk" (what
        `Ruleset.osawa_discard` is): `card.color is not None` holds for the kUnknownColor enum too, so an
        un-hinted card tests `fireworks[-1]`, the last colour's stack.  "intended": un-hinted colours are
        skipped.  The binding that decides which one the reference really executes is not available
        (SURVEY RB-3); a rule list uses one reading for all its osawa_discard entries."""
        if unknown_color not in ("quirk", "intended"):
            raise ValueError("unknown_color must be 'quirk' or 'intended'")
        return Rule(1, "osawa_discard", iarg=int(unknown_color == "intended"))

    @staticmethod
    def tell_dispensable_factory(min_information_tokens=8):         # :454
        return Rule(7, "tell_dispensable", iarg=min_information_tokens)

    @staticmethod
    def play_probably_safe_factory(treshold=0.95, require_extra_lives=False):  # :617
        return Rule(9, "play_probably_safe", farg=treshold, iarg=int(bool(require_extra_lives)))

    # The reference versions of the next three index the observation dict-style and raise TypeError
    # on the pybind objects they are given (ruleset.py:522-535, 570-594, 638-649); these follow the
    # semantics the code spells out.
    tell_anyone_useless_card = Rule(14, "tell_anyone_useless_card")  # :522
    tell_playable_card = Rule(15, "tell_playable_card")              # :570 (coin :584 = counter-based draw)

    @staticmethod
    def discard_probably_useless_factory(treshold=0.75):             # :638
        return Rule(16, "discard_probably_useless", farg=treshold)
