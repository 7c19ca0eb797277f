"""Mirror of montecarlo.card (/root/reference/src/montecarlo/card.clj:7-13)."""
from collections import namedtuple

Card = namedtuple("Card", ["suit", "rank"])          # (defrecord Card [suit rank]) card.clj:7-8

SUITS = ("hearts", "diamonds", "spades", "clubs")    # card.clj:11 order = suit ids 0..3
RANKS = tuple(range(2, 15))                           # card.clj:12

# card.clj:10-13: suit-major, rank ascending; index in this list = card id of the C ABI
COMPLETE_DECK = tuple(Card(s, r) for s in SUITS for r in RANKS)


def card_id(card):
    """index of a deck card in COMPLETE-DECK (13*suit + rank-2)"""
    return 13 * SUITS.index(card.suit) + (card.rank - 2)


def from_wire(d):
    """{"suit": "hearts", "rank": 8} (README.md:54-55, the JSON cheshire emits at server.clj:130) -> Card"""
    return Card(d["suit"], int(d["rank"]))


def to_wire(card):
    return {"suit": card.suit, "rank": card.rank}
