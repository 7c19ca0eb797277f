"""Regenerates the "predicates" table of reference_vectors.json: every predicate assert of the reference's
test/montecarlo/hand_evaluator_test.clj:6-55, transcribed form by form (the reference is Clojure and cannot be imported,
so this is a transcription, not an execution).  Cards are [suit id, rank] of COMPLETE-DECK (card.clj:10-13: suits hearts,
diamonds, spades, clubs = 0..3, ranks 2..14); `twos` etc. = (filter #(= r (:rank %)) card/COMPLETE-DECK).
usage: python tests/golden/make_predicates.py [--check]"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(HERE, "reference_vectors.json")
DECK = [[s, r] for s in range(4) for r in range(2, 15)]
twos, threes, fours = ([c for c in DECK if c[1] == r] for r in (2, 3, 4))
mix = twos[:2] + threes[:2]
ROWS = [
    (8, "constant-coll?", True, {"coll": [1, 1, 1]}), (9, "constant-coll?", False, {"coll": [1, 1, 2]}),
    (10, "constant-coll?", False, {"coll": [1, 2, 1]}), (11, "constant-coll?", False, {"coll": [2, 1, 1]}),
    (17, "straight?", True, {"cards": DECK[:5], "shuffle": True}), (18, "straight?", False, {"cards": DECK[::2][:5]}),
    (24, "n-of-a-kind?", True, {"cards": twos, "n": 4}), (25, "n-of-a-kind?", True, {"cards": twos, "n": 3}),
    (26, "n-of-a-kind?", True, {"cards": twos, "n": 2}), (27, "n-of-a-kind?", False, {"cards": mix, "n": 4}),
    (28, "n-of-a-kind?", False, {"cards": mix, "n": 3}), (29, "n-of-a-kind?", True, {"cards": mix, "n": 2, "shuffle": True}),
    (36, "full-house?", True, {"cards": twos}), (37, "full-house?", False, {"cards": twos + threes[:1]}),
    (38, "full-house?", True, {"cards": twos[:3] + threes[:2]}), (39, "full-house?", False, {"cards": twos[:2] + threes[:1] + fours[:1]}),
    (41, "full-house?", False, {"cards": twos[:2] + threes[:2] + fours[:1]}),
    (49, "two-pair?", True, {"cards": twos}), (50, "two-pair?", True, {"cards": twos + threes[:1]}),
    (51, "two-pair?", True, {"cards": twos[:3] + threes[:2]}), (52, "two-pair?", False, {"cards": twos[:2] + threes[:1] + fours[:1]}),
    (54, "two-pair?", True, {"cards": twos[:2] + threes[:2] + fours[:1]}),
]


def table():
    out = []
    for line, fn, truthy, extra in ROWS:
        e = {"line": line, "fn": fn, "truthy": truthy}
        for k in ("coll", "cards", "n", "shuffle"):
            if k in extra:
                e[k] = extra[k]
        out.append(e)
    return out


if __name__ == "__main__":
    g = json.load(open(PATH))
    if "--check" in sys.argv:
        assert g["predicates"]["asserts"] == table(), "reference_vectors.json differs from this script"
        print("predicates table matches (%d rows)" % len(table()))
    else:
        g["predicates"]["asserts"] = table()
        json.dump(g, open(PATH, "w"), indent=1)
        print("wrote %d rows" % len(table()))
