"""Second, independent restatement of the HLE engine + canonical encoder, in plain Python.

TEST INFRASTRUCTURE.  Written object-style after the upstream library's own structure
(HanabiState with a full move history, HanabiObservation with its `last_moves` window,
CanonicalObservationEncoder with per-section encoders) -- deliberately NOT sharing the
shortcuts of oracle/hanabi_oracle.c (which keeps only the most recent non-deal move, counts
instead of piles, bitmasks instead of lists).  tests/test_hanabi_oracle.py plays random
games through both and requires identical observations, masks, rewards and terminals.
Spec: SURVEY.md Appendix A ([RECALL] of deepmind/hanabi-learning-environment).
"""
CHANCE = -1
DEAL, PLAY, DISCARD, REVEAL_COLOR, REVEAL_RANK = "deal", "play", "discard", "reveal_color", "reveal_rank"


class Game:
    def __init__(self, players=2, colors=5, ranks=5, hand_size=5, max_info=8, max_life=3):
        self.players, self.colors, self.ranks = players, colors, ranks
        self.hand_size, self.max_info, self.max_life = hand_size, max_info, max_life

    def instances(self, color, rank):
        if rank == 0:
            return 3
        if rank == self.ranks - 1:
            return 1
        return 2

    @property
    def deck_size(self):
        return sum(self.instances(c, r) for c in range(self.colors) for r in range(self.ranks))

    @property
    def num_moves(self):
        return 2 * self.hand_size + (self.players - 1) * (self.colors + self.ranks)

    def get_move(self, uid):
        H, C, R, P = self.hand_size, self.colors, self.ranks, self.players
        if uid < H:
            return dict(type=DISCARD, card_index=uid)
        uid -= H
        if uid < H:
            return dict(type=PLAY, card_index=uid)
        uid -= H
        if uid < (P - 1) * C:
            return dict(type=REVEAL_COLOR, target_offset=1 + uid // C, color=uid % C)
        uid -= (P - 1) * C
        return dict(type=REVEAL_RANK, target_offset=1 + uid // R, rank=uid % R)


class Knowledge:
    def __init__(self, game):
        self.color = None
        self.rank = None
        self.color_plausible = [True] * game.colors
        self.rank_plausible = [True] * game.ranks


class State:
    def __init__(self, game, deck):
        """deck: list of (color, rank) in dealing order."""
        self.g = game
        self.deck = list(deck)
        self.hands = [[] for _ in range(game.players)]       # lists of (color, rank)
        self.knowledge = [[] for _ in range(game.players)]
        self.discard_pile = []
        self.fireworks = [0] * game.colors
        self.info = game.max_info
        self.life = game.max_life
        self.turns_to_play = game.players
        self.history = []
        self.cur_player = CHANCE
        self.next_non_chance = 0
        while self.cur_player == CHANCE:
            self._apply_deal()

    def player_to_deal(self):
        for p, h in enumerate(self.hands):
            if len(h) < self.g.hand_size:
                return p
        return -1

    def _advance(self):
        if self.deck and self.player_to_deal() >= 0:
            self.cur_player = CHANCE
        else:
            self.cur_player = self.next_non_chance
            self.next_non_chance = (self.cur_player + 1) % self.g.players

    def _apply_deal(self):
        p = self.player_to_deal()
        card = self.deck.pop(0)
        self.hands[p].append(card)
        self.knowledge[p].append(Knowledge(self.g))
        self.history.append(dict(move=dict(type=DEAL), player=CHANCE, deal_to_player=p))
        self._advance()

    def score(self):
        return 0 if self.life <= 0 else sum(self.fireworks)

    def is_terminal(self):
        return (self.life < 1 or self.score() >= self.g.colors * self.g.ranks
                or self.turns_to_play <= 0)

    def move_is_legal(self, m):
        p = self.cur_player
        if m["type"] == DISCARD:
            return self.info < self.g.max_info and m["card_index"] < len(self.hands[p])
        if m["type"] == PLAY:
            return m["card_index"] < len(self.hands[p])
        if self.info <= 0:
            return False
        target = self.hands[(p + m["target_offset"]) % self.g.players]
        if m["type"] == REVEAL_COLOR:
            return any(c == m["color"] for c, _ in target)
        return any(r == m["rank"] for _, r in target)

    def legal_mask(self, player):
        if player != self.cur_player:
            return [0] * self.g.num_moves
        return [int(self.move_is_legal(self.g.get_move(u))) for u in range(self.g.num_moves)]

    def apply_move(self, m):
        assert self.move_is_legal(m)
        if not self.deck:
            self.turns_to_play -= 1
        p = self.cur_player
        item = dict(move=m, player=p, scored=False, information_token=False, color=-1, rank=-1,
                    reveal_bitmask=0, newly_revealed_bitmask=0)
        if m["type"] == DISCARD:
            if self.info < self.g.max_info:
                self.info += 1
                item["information_token"] = True
            card = self.hands[p].pop(m["card_index"])
            self.knowledge[p].pop(m["card_index"])
            item["color"], item["rank"] = card
            self.discard_pile.append(card)
        elif m["type"] == PLAY:
            card = self.hands[p].pop(m["card_index"])
            self.knowledge[p].pop(m["card_index"])
            item["color"], item["rank"] = card
            if card[1] == self.fireworks[card[0]]:
                self.fireworks[card[0]] += 1
                item["scored"] = True
                if self.fireworks[card[0]] == self.g.ranks and self.info < self.g.max_info:
                    self.info += 1
                    item["information_token"] = True
            else:
                self.life -= 1
                self.discard_pile.append(card)
        else:
            self.info -= 1
            t = (p + m["target_offset"]) % self.g.players
            for i, (c, r) in enumerate(self.hands[t]):
                k = self.knowledge[t][i]
                if m["type"] == REVEAL_COLOR:
                    if c == m["color"]:
                        item["reveal_bitmask"] |= 1 << i
                        if k.color is None:
                            item["newly_revealed_bitmask"] |= 1 << i
                        k.color = c
                        k.color_plausible = [j == c for j in range(self.g.colors)]
                    else:
                        k.color_plausible[m["color"]] = False
                else:
                    if r == m["rank"]:
                        item["reveal_bitmask"] |= 1 << i
                        if k.rank is None:
                            item["newly_revealed_bitmask"] |= 1 << i
                        k.rank = r
                        k.rank_plausible = [j == r for j in range(self.g.ranks)]
                    else:
                        k.rank_plausible[m["rank"]] = False
        self.history.append(item)
        self._advance()

    def step(self, uid):
        """rl_env.step: apply, resolve chance nodes, reward = score delta."""
        before = self.score()
        self.apply_move(self.g.get_move(uid))
        while self.cur_player == CHANCE:
            self._apply_deal()
        return self.score() - before, self.is_terminal()


class Observation:
    """HanabiObservation: observer-relative view incl. the last_moves window."""

    def __init__(self, state, observer):
        g = state.g
        P = g.players
        self.g = g
        self.hands = [state.hands[(observer + k) % P] for k in range(P)]
        self.knowledge = [state.knowledge[(observer + k) % P] for k in range(P)]
        self.discard_pile = state.discard_pile
        self.fireworks = state.fireworks
        self.deck_size = len(state.deck)
        self.info, self.life = state.info, state.life
        # history since the observer's own last move, most recent first; initial deals excluded
        hist = state.history
        first_player_move = next((i for i, it in enumerate(hist) if it["player"] != CHANCE), len(hist))
        self.last_moves = []
        for i in range(len(hist) - 1, first_player_move - 1, -1):
            it = dict(hist[i])
            if it["move"]["type"] == DEAL:
                it["deal_to_player"] = (it["deal_to_player"] - observer + P) % P
            else:
                it["player"] = (it["player"] - observer + P) % P
            self.last_moves.append(it)
            if hist[i]["player"] == observer:
                break


def encode(obs):
    g = obs.g
    P, H, C, R = g.players, g.hand_size, g.colors, g.ranks
    bpc = C * R
    out = []

    # hands
    sec = [0] * ((P - 1) * H * bpc + P)
    off = 0
    for k in range(1, P):
        for i, (c, r) in enumerate(obs.hands[k]):
            sec[off + i * bpc + c * R + r] = 1
        off += H * bpc
    for k in range(P):
        if len(obs.hands[k]) < H:
            sec[off + k] = 1
    out += sec

    # board
    sec = [0] * (g.deck_size - P * H + C * R + g.max_info + g.max_life)
    for i in range(obs.deck_size):
        sec[i] = 1
    off = g.deck_size - P * H
    for c in range(C):
        if obs.fireworks[c] > 0:
            sec[off + obs.fireworks[c] - 1] = 1
        off += R
    for i in range(obs.info):
        sec[off + i] = 1
    off += g.max_info
    for i in range(obs.life):
        sec[off + i] = 1
    out += sec

    # discards
    sec = []
    for c in range(C):
        for r in range(R):
            n = sum(1 for card in obs.discard_pile if card == (c, r))
            inst = g.instances(c, r)
            sec += [1] * n + [0] * (inst - n)
    out += sec

    # last action
    sec = [0] * (P + 4 + P + C + R + H + H + bpc + 2)
    last = next((it for it in obs.last_moves if it["move"]["type"] != DEAL), None)
    if last is not None:
        m = last["move"]
        off = 0
        sec[off + last["player"]] = 1
        off += P
        sec[off + [PLAY, DISCARD, REVEAL_COLOR, REVEAL_RANK].index(m["type"])] = 1
        off += 4
        if m["type"] in (REVEAL_COLOR, REVEAL_RANK):
            sec[off + (last["player"] + m["target_offset"]) % P] = 1
        off += P
        if m["type"] == REVEAL_COLOR:
            sec[off + m["color"]] = 1
        off += C
        if m["type"] == REVEAL_RANK:
            sec[off + m["rank"]] = 1
        off += R
        if m["type"] in (REVEAL_COLOR, REVEAL_RANK):
            for i in range(H):
                if last["reveal_bitmask"] & (1 << i):
                    sec[off + i] = 1
        off += H
        if m["type"] in (PLAY, DISCARD):
            sec[off + m["card_index"]] = 1
        off += H
        if m["type"] in (PLAY, DISCARD):
            sec[off + last["color"] * R + last["rank"]] = 1
        off += bpc
        if m["type"] == PLAY:
            if last["scored"]:
                sec[off] = 1
            if last["information_token"]:
                sec[off + 1] = 1
    out += sec

    # card knowledge
    per = bpc + C + R
    sec = [0] * (P * H * per)
    for k in range(P):
        for i, kn in enumerate(obs.knowledge[k]):
            o = (k * H + i) * per
            for c in range(C):
                for r in range(R):
                    if kn.color_plausible[c] and kn.rank_plausible[r]:
                        sec[o + c * R + r] = 1
            if kn.color is not None:
                sec[o + bpc + kn.color] = 1
            if kn.rank is not None:
                sec[o + bpc + C + kn.rank] = 1
    out += sec
    return out
