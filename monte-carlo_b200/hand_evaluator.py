"""Mirror of montecarlo.hand-evaluator (/root/reference/src/montecarlo/hand_evaluator.clj): same var names
(with - -> _ and ? -> _p), same argument meaning, same return shapes, so tests read like the reference's
test/montecarlo/hand_evaluator_test.clj.

`evaluator`, `player_to_hand_value` (and `max_hand_value` through key order) are answered by the CUDA
library; the small predicates are test surface in the reference and stay plain host code here
exactly as they would stay plain Clojure in the JVM drop-in.

Differences from the reference, all deliberate:
  Q2  a full house's slot 1 is a list, and two full houses compare lexicographically instead of raising
      ClassCastException;
  Q3  fewer than 5 cards raises ValueError (the reference raises ArityException);
  Q7  player_to_hand_value does not print.
"""
from itertools import combinations, permutations

from . import api
from .engine import McError, MC_ERR_TOO_FEW


def _get(card, att):
    return card[att] if isinstance(card, dict) else getattr(card, att)


# ---- plain predicates (hand_evaluator.clj:6-94) -------------------------------------------------------
def constant_coll_p(coll):                                   # :6-9
    coll = list(coll)
    return coll if all(a == b for a, b in zip(coll, coll[1:])) else None


def constant_cards_p(cards, att):                            # :11-14
    cards = list(cards)
    return cards if constant_coll_p([_get(c, att) for c in cards]) is not None else None


def high_card(card_a, card_b):                               # :18-22
    return card_a if _get(card_a, "rank") > _get(card_b, "rank") else card_b


def high_hand(hand_a, hand_b):                               # :24-30
    from functools import reduce
    high_a, high_b = reduce(high_card, hand_a), reduce(high_card, hand_b)
    return hand_a if high_card(high_a, high_b) == high_a else hand_b


def straight_p(cards):                                       # :32-40 (no ace-low straight)
    cards = list(cards)
    ranks = sorted(_get(c, "rank") for c in cards)
    return cards if all(a + 1 == b for a, b in zip(ranks, ranks[1:])) else None


def flush_p(cards):                                          # :42-44
    return constant_cards_p(cards, "suit")


def n_of_a_kinds(cards, n):                                  # :46-49
    return [list(c) for c in combinations(list(cards), n) if constant_cards_p(c, "rank") is not None]


def n_of_a_kind_p(cards, n):                                 # :51-55
    from functools import reduce
    kinds = n_of_a_kinds(cards, n)
    return reduce(high_hand, kinds) if kinds else None


def four_of_a_kind_p(cards):                                 # :59-61
    return n_of_a_kind_p(cards, 4)


def three_of_a_kind_p(cards):                                # :63-65
    return n_of_a_kind_p(cards, 3)


def pair_p(cards):                                           # :67-69
    return n_of_a_kind_p(cards, 2)


def n_pairs_p(cards, k1, k2):                                # :71-79
    cards = list(cards)
    for perm in permutations(cards):
        g1 = sorted(perm[:k1], key=lambda c: _get(c, "rank"))
        g2 = sorted(perm[k1:k1 + k2], key=lambda c: _get(c, "rank"))
        if constant_cards_p(g1, "rank") is not None and constant_cards_p(g2, "rank") is not None:
            return [g1, g2]
    return None


def full_house_p(cards):                                     # :88-90
    return n_pairs_p(cards, 3, 2)


def two_pair_p(cards):                                       # :92-94
    return n_pairs_p(cards, 2, 2)


def cards_to_ranks(cards):                                   # :96-98
    return sorted((_get(c, "rank") for c in cards), reverse=True)


# ---- GPU-backed values ----------------------------------------------------------------------------------
def _pack(cards):
    """opaque suits -> ids 0..3 by first appearance, integer ranks.

    Four of the reference's own test vectors (hand_evaluator_test.clj:92,99,105,125) hold the SAME card
    twice, which a card set cannot express.  The reference only ever asks whether ALL suits are equal
    (flush? :42-44), so when a 5-card hand is not single-suited a repeated card can be moved to any free
    suit without changing the value, as long as at least two suits stay in use."""
    suits, ids, ranks = {}, [], []
    for c in cards:
        s = _get(c, "suit")
        if s not in suits:
            if len(suits) == 4:
                raise ValueError("more than four distinct suits")
            suits[s] = len(suits)
        ids.append(suits[s])
        ranks.append(int(_get(c, "rank")))
    if len(set(zip(ids, ranks))) != len(ids):
        if len(set(ids)) < 2:
            raise ValueError("repeated card in a single-suited hand: no card set expresses it")
        seen = set()
        for i, (s, r) in enumerate(zip(ids, ranks)):
            if (s, r) in seen:
                free = [t for t in range(4) if (t, r) not in seen and (t, r) not in set(zip(ids[i + 1:], ranks[i + 1:]))]
                if not free:
                    raise ValueError("more than four cards of one rank")
                ids[i] = s = free[0]
            seen.add((s, r))
        if len(set(ids)) < 2:
            raise ValueError("cannot re-suit the repeated cards")
    return ids, ranks


def _value(cards):
    cards = list(cards)
    ids, ranks = _pack(cards)
    try:
        return api.key_decode(api.eval_cards(ids, ranks))
    except McError as e:
        if e.code == MC_ERR_TOO_FEW:
            raise ValueError("fewer than 5 cards (the reference raises ArityException)") from e
        raise


def evaluator(cards):                                        # :112-133, exactly 5 cards
    cards = list(cards)
    if len(cards) != 5:
        raise ValueError("evaluator takes 5 cards")
    return _value(cards)


def max_hand_value(val_a, val_b):                            # :156-160 (tie -> val_b)
    return val_a if _compare(val_a, val_b) > 0 else val_b


def _compare(a, b):
    """clojure.core/compare on [cat hit kick]: length first, then element-wise (APersistentVector.compareTo)"""
    def cmp(x, y):
        if isinstance(x, (list, tuple)):
            if len(x) != len(y):
                return -1 if len(x) < len(y) else 1
            for p, q in zip(x, y):
                c = cmp(p, q)
                if c:
                    return c
            return 0
        return (x > y) - (x < y)
    return cmp(a, b)


def player_to_hand_value(board, player_id):                  # :162-172
    """board: object/dict with community_cards (deal order) and room; the pocket comes from
    database.player_hand as in the reference (:169).  Here `board` may also carry the pockets directly:
    board["hands"][player_id]."""
    community = list(_get(board, "community_cards"))
    pocket = list(_get(board, "hands")[player_id])
    return _value(community + pocket)
