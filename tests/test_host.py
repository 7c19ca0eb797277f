"""CPU-only: host-side logic, the C-ABI library's exports, and the product evaluator's formulation compiled
for the host (tests/host_check.cpp) against the oracle on every 5-, 6- and 7-card hand."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
Card = None


def test_library_loads_and_exports_every_declared_symbol(pkg):
    L = pkg.engine.lib()
    header = open(os.path.join(ROOT, "include", "mcpoker.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(mc_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    assert declared == set(pkg.engine.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert L.mc_abi_version() == 2


def test_header_is_plain_c_and_links(pkg, tmp_path):
    """include/mcpoker.h compiles as strict C99 and a C program linked against libmcpoker.so (what a JNA / Panama /
    cgo binding relies on) runs the device-free part of the ABI"""
    import subprocess
    exe = str(tmp_path / "abi_consumer")
    libdir = os.path.dirname(pkg.engine.LIB_PATH)
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c", "abi_consumer.c"), "-o", exe,
                           "-L", libdir, "-lmcpoker", "-Wl,-rpath," + libdir])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, (out.returncode, out.stdout, out.stderr)
    assert "abi consumer ok" in out.stdout


def test_pack48_wire_format(pkg, oracle):
    """api.pack48: 6 bytes per hand, seven 6-bit card positions lowest card first, unused fields 0; decodes back"""
    import numpy as np
    for n_cards in (5, 6, 7):
        masks = oracle.random_hands(3, 0, 5000, n_cards)
        p = pkg.api.pack48(masks)
        assert p.dtype == np.uint8 and p.shape == (6 * 5000,)
        b = p.reshape(-1, 6).astype(np.uint64)
        h = sum(b[:, k] << np.uint64(8 * k) for k in range(6))
        back = np.zeros(5000, dtype=np.uint64)
        for k in range(7):
            f = (h >> np.uint64(6 * k)) & np.uint64(63)
            assert np.all((f != 0) == (k < n_cards))
            back |= np.where(f != 0, np.uint64(1) << f, np.uint64(0))
        assert np.array_equal(back, masks)
    first = int(masks[0])
    lowest = (first & -first).bit_length() - 1
    assert int(p[0]) & 63 == lowest
    with pytest.raises(ValueError):
        pkg.api.pack48(np.array([0xFF00FF], dtype=np.uint64))


def test_colex_rank_wire_format(pkg, oracle):
    """api.colex_rank (what a caller of mc_eval_batch_colex32 computes): rank of the i-th hand of the oracle's colex
    enumeration is i; random hands round-trip through the oracle's enumeration"""
    import numpy as np
    for first in (0, 1_000_000, 133784560 - 5000):
        masks = oracle.enumerate_masks(7, first, 5000)
        assert np.array_equal(pkg.api.colex_rank(masks), np.arange(first, first + 5000, dtype=np.uint32))
    rnd = oracle.random_hands(7, 0, 2000)
    ranks = pkg.api.colex_rank(rnd)
    back = np.array([oracle.enumerate_masks(7, int(r), 1)[0] for r in ranks[:200]], dtype=np.uint64)
    assert np.array_equal(back, rnd[:200])
    with pytest.raises(ValueError):
        pkg.api.colex_rank(oracle.random_hands(7, 0, 10, 6))


def test_pack27_wire_format(pkg):
    """api.pack27: 27 bits per hand, little-endian bit stream; every field reads back, ragged sizes"""
    import numpy as np
    rng = np.random.default_rng(3)
    for n in (1, 31, 32, 33, 1000, 4097):
        r = rng.integers(0, 133784560, n, dtype=np.uint32)
        s = pkg.api.pack27(r)
        assert s.dtype == np.uint32 and s.shape[0] == (27 * n + 31) // 32
        bits = int.from_bytes(s.tobytes(), "little")
        back = [(bits >> (27 * j)) & 0x7FFFFFF for j in range(n)]
        assert back == r.tolist()
    with pytest.raises(ValueError):
        pkg.api.pack27(np.array([1 << 27], dtype=np.uint32))


def test_no_gpu_means_loud_failure_not_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.engine.McError) as e:
        pkg.engine.init()
    assert e.value.code == pkg.engine.MC_ERR_NO_DEVICE
    with pytest.raises(pkg.engine.McError) as e:
        pkg.engine.check(pkg.engine.lib().mc_enumerate7((C.c_uint64 * 9)(), None))
    assert e.value.code == pkg.engine.MC_ERR_NOT_INIT


def test_pure_helpers(pkg, oracle):
    api = pkg.api
    assert [api.card_bit(i) for i in (0, 12, 13, 51)] == [oracle.card_bit(i) for i in (0, 12, 13, 51)]
    assert api.card_bit(52) == 0
    assert api.key_decode(0x8_98765) == [8, [9, 8, 7, 6, 5], []]
    assert api.key_decode(0x7_11113) == [7, [1, 1, 1, 1], [3]]
    assert api.key_decode(0x6_88899) == [6, [8, 8, 8, 9, 9], []]
    assert api.key_decode(0x3_55582) == [3, [5, 5, 5], [8, 2]]
    assert api.key_decode(0x2_55228) == [2, [5, 5, 2, 2], [8]]
    assert api.key_decode(0x1_55876) == [1, [5, 5], [8, 7, 6]]
    assert api.key_decode(0x0_87642) == [0, [8, 7, 6, 4, 2], []]
    with pytest.raises(pkg.engine.McError):
        api.key_decode(0xFFFFFFFF)
    # packed <-> nibble: cat 1, pair of 5s (bit 4), kickers 8 7 6 (bits 7,6,5)
    packed = (1 << 28) | ((1 << 4) << 14) | (1 << 7 | 1 << 6 | 1 << 5)
    assert api.key_packed_to_nibble(packed) == 0x155876
    assert api.hole_class(14, 14, False) == 168 and api.hole_class(13, 13, False) == 154
    assert api.hole_class(14, 13, True) == 13 * 12 + 11 and api.hole_class(14, 13, False) == 13 * 11 + 12


def test_card_module_matches_card_clj(pkg):
    card = pkg.card
    assert len(card.COMPLETE_DECK) == 52
    assert card.COMPLETE_DECK[0] == card.Card("hearts", 2) and card.COMPLETE_DECK[51] == card.Card("clubs", 14)
    assert [card.card_id(c) for c in card.COMPLETE_DECK] == list(range(52))
    assert card.from_wire({"suit": "hearts", "rank": 8}) == card.Card("hearts", 8)


def test_mirror_predicates_like_the_reference_tests(pkg):
    """hand_evaluator_test.clj:6-55 against the plain host predicates of the mirror module"""
    import random
    he, card = pkg.hand_evaluator, pkg.card
    deck = list(card.COMPLETE_DECK)
    assert he.constant_coll_p([1, 1, 1])
    assert not he.constant_coll_p([1, 1, 2]) and not he.constant_coll_p([1, 2, 1]) and not he.constant_coll_p([2, 1, 1])
    h1 = deck[:5]
    random.Random(0).shuffle(h1)
    assert he.straight_p(h1) and not he.straight_p(deck[::2][:5])
    twos = [c for c in deck if c.rank == 2]
    threes = [c for c in deck if c.rank == 3]
    fours = [c for c in deck if c.rank == 4]
    assert he.n_of_a_kind_p(twos, 4) and he.n_of_a_kind_p(twos, 3) and he.n_of_a_kind_p(twos, 2)
    assert not he.n_of_a_kind_p(twos[:2] + threes[:2], 4) and not he.n_of_a_kind_p(twos[:2] + threes[:2], 3)
    assert he.n_of_a_kind_p(twos[:2] + threes[:2], 2)
    assert he.full_house_p(twos) and not he.full_house_p(twos + threes[:1])
    assert he.full_house_p(twos[:3] + threes[:2])
    assert not he.full_house_p(twos[:2] + threes[:1] + fours[:1])
    assert not he.full_house_p(twos[:2] + threes[:2] + fours[:1])
    assert he.two_pair_p(twos) and he.two_pair_p(twos + threes[:1]) and he.two_pair_p(twos[:3] + threes[:2])
    assert not he.two_pair_p(twos[:2] + threes[:1] + fours[:1])
    assert he.two_pair_p(twos[:2] + threes[:2] + fours[:1])
    assert he.max_hand_value([1, [5, 5], [8, 7, 6]], [0, [8, 7, 6, 4, 2], []]) == [1, [5, 5], [8, 7, 6]]
    assert he.max_hand_value([6, [8, 8, 8, 9, 9], []], [6, [9, 9, 9, 8, 8], []]) == [6, [9, 9, 9, 8, 8], []]


def test_mirror_predicate_trace_table(pkg):
    """every predicate assert of hand_evaluator_test.clj:6-55 (tests/golden/reference_vectors.json "predicates", the same
    table INTEGRATION.md traces through the Clojure drop-in) against the Python mirror: the two mirrors implement
    n-pairs? by the same permutation search, so they cannot disagree on an assert without this test failing"""
    import json
    import random
    he, card = pkg.hand_evaluator, pkg.card
    table = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_vectors.json")))["predicates"]["asserts"]
    assert len(table) == 22
    fns = {"constant-coll?": lambda e, cs: he.constant_coll_p(e["coll"]), "straight?": lambda e, cs: he.straight_p(cs),
           "n-of-a-kind?": lambda e, cs: he.n_of_a_kind_p(cs, e["n"]), "full-house?": lambda e, cs: he.full_house_p(cs),
           "two-pair?": lambda e, cs: he.two_pair_p(cs)}
    for e in table:
        cs = [card.COMPLETE_DECK[13 * s + r - 2] for s, r in e.get("cards", [])]
        orders = [cs]
        if e.get("shuffle"):
            rng = random.Random(e["line"])
            for _ in range(10):
                c2 = cs[:]
                rng.shuffle(c2)
                orders.append(c2)
        for o in orders:
            got = fns[e["fn"]](e, o)
            assert (got is not None) == e["truthy"], (e["line"], e["fn"])
    # the shape n-pairs? returns where the reference's own test relies on a SHORT second group (:36, :49)
    twos = [c for c in card.COMPLETE_DECK if c.rank == 2]
    g1, g2 = he.full_house_p(twos)
    assert len(g1) == 3 and len(g2) == 1
    g1, g2 = he.two_pair_p(twos)
    assert len(g1) == 2 and len(g2) == 2


def test_pack_resuits_repeated_cards(pkg):
    he, card = pkg.hand_evaluator, pkg.card
    C_ = card.Card
    ids, ranks = he._pack([C_("h", 5), C_("d", 2), C_("s", 3), C_("c", 5), C_("h", 5)])   # test.clj:92
    assert ranks == [5, 2, 3, 5, 5] and len(set(zip(ids, ranks))) == 5 and len(set(ids)) >= 2
    with pytest.raises(ValueError):
        he._pack([C_("h", 5), C_("h", 5), C_("h", 2), C_("h", 3), C_("h", 4)])


def test_group_winners_and_update_stacks(pkg):
    gp = pkg.gameplay
    entries = [{"player": "a", "hand_value": [3, [5, 5, 5], [8, 2]]}, {"player": "b", "hand_value": [3, [5, 5, 5], [8, 2]]},
               {"player": "c", "hand_value": [1, [5, 5], [8, 7, 6]]}]
    assert gp.group_winners(entries) == [{"a", "b"}, {"c"}]
    stacks = gp.update_stacks([{"bet": 10, "n": 3, "players": {"a", "b", "c"}}, {"bet": 5, "n": 1, "players": {"c"}}],
                              [{"a", "b"}, {"c"}], {"a": 0, "b": 0, "c": 0})
    assert stacks == {"a": 15, "b": 15, "c": 5}


def test_product_evaluator_formulation_on_every_hand(oracle):
    """eval.cuh (the product's device evaluator, plain integer code) compiled with g++ == oracle on all
    C(52,5)+C(52,6)+C(52,7) hands and on the 56-card deck that includes the tests' rank 1."""
    exe = "/tmp/mcp_host_check"
    subprocess.check_call(["/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++", "-O2", "-fopenmp", "-o", exe,
                           os.path.join(ROOT, "tests", "host_check.cpp"), "-L" + oracle.ORACLE_DIR, "-loracle",
                           "-Wl,-rpath," + oracle.ORACLE_DIR])
    for arg in ("0", "1", "2"):       # 2 = the additive-key / perfect-hash evaluator (hashtab.cuh) on all C(52,7) hands
        out = subprocess.run([exe, arg], capture_output=True, text=True)
        assert out.returncode == 0, out.stdout + out.stderr
        assert "mismatches=0" in out.stdout


def test_settle_pots_matches_oracle_and_mirror(pkg, oracle):
    """mc_settle_pots (update-stacks, gameplay.clj:104-116) == the oracle's literal restatement == the Python mirror,
    on random side-pot structures; pure host integer arithmetic, so it runs without a GPU."""
    import random
    rnd = random.Random(11)
    for _ in range(300):
        n = rnd.randint(2, 9)
        n_groups = rnd.randint(1, n)
        group_of = [rnd.randrange(n_groups) for _ in range(n)]
        used = sorted(set(group_of))
        group_of = [used.index(g) for g in group_of]                     # groups 0..k-1, all non-empty
        pots = []
        for _ in range(rnd.randint(0, 4)):
            players = rnd.sample(range(n), rnd.randint(1, n))
            pots.append((rnd.randint(1, 500), rnd.randint(1, n), players))
        got = pkg.api.settle_pots(group_of, pots)
        sets = [sum(1 << s for s in range(n) if group_of[s] == g) for g in range(len(used))]
        stack = (C.c_longlong * 32)()
        rc = oracle.lib().orc_update_stacks((C.c_uint * len(sets))(*sets), len(sets),
                                            (C.c_longlong * max(1, len(pots)))(*[p[0] for p in pots]),
                                            (C.c_int * max(1, len(pots)))(*[p[1] for p in pots]),
                                            (C.c_uint * max(1, len(pots)))(*[sum(1 << s for s in p[2]) for p in pots]),
                                            len(pots), stack)
        assert rc == 0 and got == list(stack[:n])
        winners = [{s for s in range(n) if group_of[s] == g} for g in range(len(used))]
        mirror = pkg.gameplay.update_stacks([{"bet": b, "n": k, "players": set(pl)} for b, k, pl in pots], winners, {})
        assert got == [mirror.get(s, 0) for s in range(n)]
    # a pot whose players all folded: the reference divides by zero; the ABI reports MC_ERR_INVALID
    with pytest.raises(pkg.engine.McError):
        pkg.api.settle_pots([0, -1], [(10, 2, [1])])


def test_canonical_boards_cover_every_board(pkg):
    """the suit-canonical 5-card boards of the per-board matrix: 134,459 orbits whose sizes sum to
    C(52,5); host-only enumeration, no GPU"""
    n, w = pkg.api.preflop_boards_count()
    assert n == 134459 and w == 2598960
