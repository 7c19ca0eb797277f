"""GPU parity tests (run with -m gpu on a B200): every entry point of the C ABI against the CPU oracle on
the same inputs -- bit-exact, because the whole path is integer work."""
import json
import os
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_vectors.json")))
SG = GOLD["survey_goldens"]
SEED = 0x5EED5EED


def test_native_library_is_loaded(gpu):
    assert gpu.engine.lib().mc_device_count() >= 1
    maps = open("/proc/self/maps").read()
    assert "libmcpoker.so" in maps


# ---- the reference's own known-answer vectors, through the mirror of its API -------------------------
@pytest.mark.parametrize("vec", GOLD["evaluator"], ids=lambda v: "line%d" % v["line"])
def test_evaluator_reference_vectors(gpu, vec):              # hand_evaluator_test.clj:57-137
    he, Card = gpu.hand_evaluator, gpu.card.Card
    suit = {0: "h", 1: "d", 2: "s", 3: "c"}                   # the test file's keywords :h :d :s :c
    cards = [Card(suit[s], r) for s, r in vec["cards"]]
    assert he.evaluator(cards) == vec["want"]
    if vec["shuffle"]:
        rng = random.Random(vec["line"])
        for _ in range(5):
            rng.shuffle(cards)
            assert he.evaluator(cards) == vec["want"]


def test_evaluator_errors(gpu):
    he, Card = gpu.hand_evaluator, gpu.card.Card
    with pytest.raises(ValueError):                           # < 5 cards
        he._value([Card("h", 2), Card("d", 3)])
    with pytest.raises(gpu.engine.McError):
        gpu.api.eval_cards([0, 0, 0, 0, 0], [2, 3, 4, 5, 15])
    with pytest.raises(gpu.engine.McError):
        gpu.api.eval_cards([0, 0, 0, 0, 4], [2, 3, 4, 5, 6])
    # no wheel
    assert he.evaluator([Card("h", 14), Card("d", 2), Card("s", 3), Card("c", 4), Card("h", 5)]) == [0, [14, 5, 4, 3, 2], []]


def test_player_to_hand_value_5_6_7_cards(gpu, oracle):       # flop / turn / river, SURVEY Q3
    he, deck = gpu.hand_evaluator, list(gpu.card.COMPLETE_DECK)
    rng = random.Random(7)
    for n_board in (3, 4, 5):
        for _ in range(10):
            cs = rng.sample(deck, n_board + 2)
            board = {"community_cards": cs[:n_board], "hands": {"p": cs[n_board:]}}
            got = he.player_to_hand_value(board, "p")
            m = oracle.mask_of([(gpu.card.SUITS.index(c.suit), c.rank) for c in cs])
            assert got == gpu.api.key_decode(oracle.lib().orc_eval_mask(m))


# ---- K1: batch evaluation ------------------------------------------------------------------------
def test_random_hand_generator_matches_oracle(gpu, oracle):
    n = 200_000
    assert np.array_equal(gpu.api.random_hands(SEED, 0, n), oracle.random_hands(SEED, 0, n))
    assert np.array_equal(gpu.api.random_hands(SEED, 12345, 1000, 5), oracle.random_hands(SEED, 12345, 1000, 5))


@pytest.mark.parametrize("fmt", [0, 1, 2, 3, 4], ids=["nibble_auto", "packed", "nibble_flo", "nibble_rankset_table", "nibble_hash"])
def test_eval7_batch_config1_one_million_hands(gpu, oracle, fmt):
    """BASELINE config 1: 1,000,000 Philox-seeded 7-card hands, 100% equal to the oracle"""
    masks = oracle.random_hands(SEED, 0, 1_000_000)
    want = oracle.eval_batch(masks, 0)
    got = gpu.api.eval7_batch(masks, key_format=fmt)
    if fmt == 1:
        got = np.array([gpu.api.key_packed_to_nibble(int(k)) for k in got[:20000]], dtype=np.uint32)
        want = want[:20000]
        packed = gpu.api.eval7_batch(masks, key_format=1)
        # packed keys order exactly like nibble keys
        o1 = np.argsort(packed[:20000], kind="stable")
        assert np.all(np.diff(want[o1].astype(np.int64)) >= 0)
    assert np.array_equal(got, want)


def test_eval7_batch_edge_cases(gpu, oracle):
    api = gpu.api
    assert api.eval7_batch(np.zeros(0, dtype=np.uint64)).shape == (0,)
    # ragged sizes around the vector width, unaligned views, 5/6-card hands, rank 1, invalid masks
    base = oracle.random_hands(SEED, 0, 4099)
    for n in (1, 2, 3, 4, 5, 7, 8, 511, 1025, 4099):
        assert np.array_equal(api.eval7_batch(base[:n]), oracle.eval_batch(base[:n], 0))
    assert np.array_equal(api.eval7_batch(base[1:1000]), oracle.eval_batch(base[1:1000], 0))
    mixed = np.concatenate([oracle.random_hands(SEED, 0, 1000, 5), oracle.random_hands(SEED, 0, 1000, 6)])
    assert np.array_equal(api.eval7_batch(mixed), oracle.eval_batch(mixed, 0))
    rank1 = np.array([oracle.mask_of([(0, 1), (1, 2), (2, 3), (3, 5), (0, 4)]),
                      oracle.mask_of([(0, 1), (1, 1), (2, 1), (3, 1), (0, 3)])], dtype=np.uint64)
    assert [api.key_decode(k) for k in api.eval7_batch(rank1)] == [[4, [5, 4, 3, 2, 1], []], [7, [1, 1, 1, 1], [3]]]
    bad = np.array([0, 0b111100, 1 << 0 | 0x7C, 1 << 15 | 0x7C, (1 << 64) - 1, 0xFF << 2], dtype=np.uint64)
    got = api.eval7_batch(bad)
    assert np.all(got == 0xFFFFFFFF)
    assert np.array_equal(got, oracle.eval_batch(bad, 0))


def test_eval7_batch_hash_kernel_slow_path_and_ragged(gpu, oracle):
    """K1 hash kernel (key_format 4): 5/6-card hands, the tests' rank 1 and invalid masks leave the table path; odd sizes"""
    api = gpu.api
    base = oracle.random_hands(SEED + 7, 0, 50_001)
    mixed = np.concatenate([base, oracle.random_hands(SEED, 0, 3000, 5), oracle.random_hands(SEED, 0, 3000, 6),
                            np.array([oracle.mask_of([(0, 1), (1, 2), (2, 3), (3, 5), (0, 4)]),
                                      oracle.mask_of([(0, 1), (1, 1), (2, 1), (3, 1), (0, 3)]),
                                      0, 0b111100, 1 << 0 | 0x7C, 1 << 15 | 0x7C, (1 << 64) - 1, 0xFF << 2,
                                      0x7F << 2 | 1 << 15, 0x3F << 2 | 1 << 17], dtype=np.uint64)])
    rng = np.random.default_rng(5)
    rng.shuffle(mixed)
    for n in (len(mixed), 4097, 33, 2, 1):
        assert np.array_equal(api.eval7_batch(mixed[:n], key_format=4), oracle.eval_batch(mixed[:n], 0))


def test_eval7_batch_multi_stage_pipeline(gpu, oracle):
    """more hands than one staging buffer holds (4 Mi): exercises the 3-slot H2D/kernel/D2H pipeline"""
    n = (1 << 22) * 2 + 12345
    masks = oracle.random_hands(SEED + 1, 0, n)
    assert np.array_equal(gpu.api.eval7_batch(masks), oracle.eval_batch(masks, 0))


def test_k1_on_structured_hand_families(gpu, oracle):
    """BASELINE target: 100% rank agreement on every seven-card hand.  Key-by-key here for six families of
    C(49,4) = 211,876 hands (all hands containing a fixed 3-card prefix); every one of the 133,784,560 hands
    is covered through the histogram + key checksum of test_enumerate7_goldens (same device evaluator)."""
    import itertools
    bits = np.array([oracle.card_bit(i) for i in range(52)], dtype=np.uint64)
    # all hands that contain a fixed 3-card prefix: C(49,4) = 211,876 each; prefixes spread over the deck
    for prefix in [(0, 1, 2), (0, 13, 26), (12, 25, 38), (5, 18, 44), (47, 49, 51), (3, 4, 5)]:
        rest = [i for i in range(52) if i not in prefix]
        ids = np.array(list(itertools.combinations(rest, 4)), dtype=np.int64)
        masks = np.bitwise_or.reduce(bits[ids], axis=1) | np.bitwise_or.reduce(bits[list(prefix)])
        assert np.array_equal(gpu.api.eval7_batch(masks), oracle.eval_batch(masks, 0))
        assert np.array_equal(gpu.api.eval7_batch(masks, key_format=4), oracle.eval_batch(masks, 0))


def test_eval7_batch_from_library_pinned_buffers(gpu, oracle):
    """mc_alloc_host / mc_free_host: the host-buffer call on page-locked buffers handed out by the library"""
    import ctypes as C
    L = gpu.engine.lib()
    n = 300001
    masks = oracle.random_hands(SEED + 9, 0, n)
    pm, pk = L.mc_alloc_host(8 * n), L.mc_alloc_host(4 * n)
    assert pm and pk
    C.memmove(pm, masks.ctypes.data, 8 * n)
    gpu.engine.check(L.mc_eval7_batch(C.cast(pm, C.POINTER(C.c_uint64)), n, C.cast(pk, C.POINTER(C.c_uint32))))
    got = np.ctypeslib.as_array(C.cast(pk, C.POINTER(C.c_uint32)), shape=(n,)).copy()
    assert np.array_equal(got, oracle.eval_batch(masks, 0))
    assert L.mc_free_host(pm) == 0 and L.mc_free_host(pk) == 0 and L.mc_free_host(None) == 0


def test_eval_batch_packed48_equals_mask_path_and_oracle(gpu, oracle):
    """the 6-byte wire format (seven 6-bit card positions) gives the keys of the mask path: 7-, 6- and 5-card hands,
    ragged sizes around the 4-hand groups of the expander and across pipeline stages"""
    api = gpu.api
    for n_cards, n in ((7, 1_000_003), (6, 40_001), (5, 40_002), (7, 1), (7, 2), (7, 3), (7, 5), (7, (1 << 22) + 7)):
        masks = oracle.random_hands(SEED + n_cards, 11, n, n_cards)
        keys = api.eval_batch_packed48(api.pack48(masks))
        assert np.array_equal(keys, oracle.eval_batch(masks, 0)), (n_cards, n)
    # the ace-low hands of the reference's own tests (rank 1) travel too
    m = np.array([sum(1 << (16 * s + r) for s, r in zip((0, 1, 2, 3, 0), (1, 2, 3, 4, 5)))], dtype=np.uint64)
    assert np.array_equal(api.eval_batch_packed48(api.pack48(m)), oracle.eval_batch(m, 0))


def test_eval_batch_packed48_rejects_malformed_hands(gpu):
    api = gpu.api

    def pk(fields, high=0):
        h = sum(f << (6 * k) for k, f in enumerate(fields)) | high << 42
        return np.frombuffer(int(h).to_bytes(8, "little")[:6], dtype=np.uint8)
    good = [2, 3, 4, 5, 6, 18, 19]
    cases = [(pk(good), False), (pk(good[:5] + [0, 0]), False), (pk([2, 2, 4, 5, 6, 18, 19]), True),
             (pk([0] * 7), True), (pk(good[:4] + [0] * 3), True), (pk(good, high=1), True),
             (pk([16, 3, 4, 5, 6, 18, 19]), True), (pk([15, 3, 4, 5, 6, 18, 19]), True)]   # bits 0 / 15 of a lane are not cards
    keys = api.eval_batch_packed48(np.concatenate([c for c, _ in cases]))
    for k, (_, bad) in zip(keys, cases):
        assert (int(k) == 0xFFFFFFFF) == bad


def test_k1_every_seven_card_hand_key_by_key(gpu, oracle):
    """BASELINE target: 100 % bit-exact rank agreement on all 133,784,560 seven-card hands, key by key, through the
    host-buffer ABI calls against the oracle's direct evaluator -- once as masks in / nibble keys out (mc_eval7_batch)
    and once in the compact wire format, colex ranks in / class ids out (mc_eval_batch_colex32: ranks first..first+n-1
    ARE the hands of the oracle's colex enumeration)"""
    total, step, bad, bad_c = 133784560, 1 << 24, 0, 0
    class_keys = gpu.api.class_keys()
    for first in range(0, total, step):
        n = min(step, total - first)
        masks = oracle.enumerate_masks(7, first, n)
        want = oracle.eval_batch(masks, 0)
        bad += int(np.count_nonzero(gpu.api.eval7_batch(masks) != want))
        ranks = np.arange(first, first + n, dtype=np.uint32)
        cls = gpu.api.eval_batch_colex32(ranks)
        bad_c += int(np.count_nonzero(class_keys[cls] != want))
        bad_c += int(np.count_nonzero(gpu.api.eval_batch_colex27(gpu.api.pack27(ranks), n) != cls))      # the 27-bit stream
    assert bad == 0 and bad_c == 0


def test_class_keys_are_the_five_card_values(gpu, oracle):
    """mc_class_keys: the ascending list of all 7,462 distinct five-card values, so class ids order like keys"""
    ck = gpu.api.class_keys()
    assert ck.shape == (7462,) and np.all(np.diff(ck.astype(np.int64)) > 0)
    five = oracle.eval_batch(oracle.enumerate_masks(5, 0, 2598960), 0)
    assert np.array_equal(ck, np.unique(five))
    with pytest.raises(gpu.engine.McError):
        gpu.engine.check(gpu.engine.lib().mc_class_keys(ck.ctypes.data, 4892))


def test_flop_and_turn_hands_every_hand_key_by_key(gpu, oracle):
    """player->hand-value on 5 and 6 cards (hand_evaluator.clj:171 at flop / turn): all C(52,5) and all
    C(52,6) hands through the flop / turn table image of K1, as keys and as class ids, against the oracle"""
    E, api = gpu.engine, gpu.api
    ck = api.class_keys()
    for k, total in ((5, 2598960), (6, 20358520)):
        masks = oracle.enumerate_masks(k, 0, total)
        want = oracle.eval_batch(masks, 0)
        assert np.array_equal(api.eval_batch(masks, E.IN_MASK64, E.KEY_NIBBLE, k), want)
        assert np.array_equal(api.eval7_batch(masks), want)                      # stage decided from the first hand
        cls = api.eval_batch(masks, E.IN_MASK64, E.KEY_CLASS16, 0)
        assert np.array_equal(ck[cls], want)
        part = masks[:1_000_001]
        assert np.array_equal(api.eval_batch(api.pack48(part), E.IN_PACKED48, E.KEY_CLASS16, 0), cls[:1_000_001])


def test_eval_batch_every_format_and_stage_hint(gpu, oracle):
    """mc_eval_batch: {masks, packed48, colex32} x {nibble keys, class ids} x stage hint {0, 5, 6, 7}; a wrong hint or a
    mixed batch only changes the path inside the kernel (slow path), never the result; ragged sizes; invalid hands"""
    E, api = gpu.engine, gpu.api
    ck = api.class_keys()
    h7 = oracle.random_hands(SEED + 21, 0, 300_001)
    mixed = np.concatenate([h7[:50_000], oracle.random_hands(SEED, 5, 50_001, 5), oracle.random_hands(SEED, 9, 50_002, 6)])
    np.random.default_rng(1).shuffle(mixed)
    for hands in (h7, mixed):
        want = oracle.eval_batch(hands, 0)
        for hint in (0, 5, 6, 7):
            assert np.array_equal(api.eval_batch(hands, E.IN_MASK64, E.KEY_NIBBLE, hint), want)
            assert np.array_equal(ck[api.eval_batch(hands, E.IN_MASK64, E.KEY_CLASS16, hint)], want)
            assert np.array_equal(ck[api.eval_batch(api.pack48(hands), E.IN_PACKED48, E.KEY_CLASS16, hint)], want)
    want7 = oracle.eval_batch(h7, 0)
    ranks = api.colex_rank(h7)
    for n in (len(h7), 4097, 33, 3, 2, 1):
        assert np.array_equal(ck[api.eval_batch_colex32(ranks[:n])], want7[:n])
        assert np.array_equal(api.eval_batch(ranks[:n], E.IN_COLEX32, E.KEY_NIBBLE, 7), want7[:n])
        assert np.array_equal(ck[api.eval_batch_colex27(api.pack27(ranks[:n]), n)], want7[:n])
    assert api.eval_batch_colex32(np.zeros(0, dtype=np.uint32)).shape == (0,)
    big = np.concatenate([ranks] * 30)                                  # 9 M hands: several pipeline stages of the bit stream
    assert np.array_equal(api.eval_batch_colex27(api.pack27(big), len(big)), api.eval_batch_colex32(big))
    # ranks past the last combination, invalid masks, the tests' rank-1 hands (no class): MC_CLASS_INVALID / MC_KEY_INVALID
    bad = np.array([133784560, 133784561, 0xFFFFFFFF, 1 << 27], dtype=np.uint32)
    assert np.all(api.eval_batch_colex32(bad) == 0xFFFF)
    assert np.all(api.eval_batch(bad, E.IN_COLEX32, E.KEY_NIBBLE, 7) == 0xFFFFFFFF)
    assert int(api.eval_batch_colex32(np.array([133784559], dtype=np.uint32))[0]) != 0xFFFF
    odd = np.array([0, 0b111100, (1 << 64) - 1, oracle.mask_of([(0, 1), (1, 2), (2, 3), (3, 5), (0, 4)])], dtype=np.uint64)
    assert np.all(api.eval_batch(odd, E.IN_MASK64, E.KEY_CLASS16, 0) == 0xFFFF)
    with pytest.raises(E.McError):
        api.eval_batch(h7[:4], 7, E.KEY_NIBBLE, 0)
    with pytest.raises(E.McError):
        api.eval_batch(h7[:4], E.IN_MASK64, 9, 0)


# ---- K2: exhaustive enumeration (BASELINE config 2) ---------------------------------------------------
def test_enumerate7_goldens(gpu):
    hist, key_sum = gpu.api.enumerate7()
    assert hist[::-1] == SG["hist7_cat8to0"]
    assert key_sum == SG["keysum7"]
    assert sum(hist) == 133784560
    hist2, none = gpu.api.enumerate7(key_sum=False)          # the fast variant (packed keys only)
    assert hist2 == hist and none is None


def test_enumerate7_ranges_match_oracle_and_add_up(gpu, oracle):
    total = 133784560
    cuts = [0, 1, 63, 64, 65, 1_000_003, 77_777_777, total - 5, total]
    acc, acc_sum = [0] * 9, 0
    for lo, hi in zip(cuts, cuts[1:]):
        h, s = gpu.api.enumerate7(lo, hi - lo)
        if hi - lo <= 2_000_000:
            ho, so, _ = oracle.enumerate_hands(7, 0, first=lo, count=hi - lo)
            assert h == ho and s == so
        acc = [a + b for a, b in zip(acc, h)]
        acc_sum += s
    assert acc[::-1] == SG["hist7_cat8to0"] and acc_sum == SG["keysum7"]
    assert gpu.api.enumerate7(5, 0) == ([0] * 9, 0)
    with pytest.raises(gpu.engine.McError):
        gpu.api.enumerate7(total, 1)


# ---- K3: exhaustive equity (BASELINE config 3) --------------------------------------------------------
@pytest.mark.parametrize("name", ["aa_kk_no_shared", "aa_kk_both_shared", "aa_kk_one_shared"])
def test_equity_exhaustive_aa_vs_kk(gpu, name):
    g = SG[name]
    holes = [gpu.api.mask_of_ids(h) for h in g["holes_ids"]]
    w, t, l, nb = gpu.api.equity_exhaustive(holes)
    assert nb == 1712304
    assert (w[0], w[1], t[0], t[1]) == (g["win_a"], g["win_b"], g["tie"], g["tie"])
    assert l == [w[1], w[0]]


def test_equity_exhaustive_multiway_partial_boards_dead_cards(gpu, oracle):
    rng = random.Random(11)
    for n_seats, n_board, n_dead in [(2, 3, 0), (3, 3, 2), (9, 3, 0), (4, 4, 1), (6, 5, 0), (3, 2, 0), (2, 1, 3), (5, 4, 0)]:
        ids = rng.sample(range(52), 2 * n_seats + n_board + n_dead)
        holes = [gpu.api.mask_of_ids(ids[2 * s:2 * s + 2]) for s in range(n_seats)]
        board = gpu.api.mask_of_ids(ids[2 * n_seats:2 * n_seats + n_board])
        dead = gpu.api.mask_of_ids(ids[2 * n_seats + n_board:])
        w, t, l, nb = gpu.api.equity_exhaustive(holes, board, dead)
        rc, wo, to, lo, nbo = oracle.equity_exhaustive(holes, board, dead)
        assert rc == 0 and (w, t, l, nb) == (wo, to, lo, nbo)
        # board ranges add up
        if nb > 10:
            cut = nb // 3
            w1, t1, l1, _ = gpu.api.equity_exhaustive(holes, board, dead, 0, cut)
            w2, t2, l2, _ = gpu.api.equity_exhaustive(holes, board, dead, cut, nb - cut)
            assert [a + b for a, b in zip(w1, w2)] == w and [a + b for a, b in zip(t1, t2)] == t


def test_equity_argument_errors(gpu):
    api, E = gpu.api, gpu.engine
    aa = api.mask_of_ids([12, 25])
    for holes, board, dead in [([aa, aa], 0, 0), ([aa, 0], 0, 0), ([aa, api.mask_of_ids([0])], 0, 0),
                               ([aa], api.mask_of_ids([12]), 0), ([aa], 1, 0), ([aa], api.mask_of_ids(range(6)), 0)]:
        with pytest.raises(E.McError) as e:
            api.equity_exhaustive(holes, board, dead)
        assert e.value.code == E.MC_ERR_INVALID


# ---- K4: Monte Carlo equity (BASELINE config 4) -------------------------------------------------------
def test_equity_mc_bit_exact_vs_oracle_9_seats(gpu, oracle):
    holes = [gpu.api.mask_of_ids([12, 25])] + [0] * 8
    n = 1 << 20
    got = gpu.api.equity_mc(holes, 0, 0, SEED, 0, n)
    rc, w, t, l = oracle.equity_mc(holes, 0, 0, SEED, 0, n)
    assert rc == 0 and got == (w, t, l)
    assert all(a + b + c == n for a, b, c in zip(*got))


def test_equity_mc_split_invariance_and_offsets(gpu, oracle):
    holes = [gpu.api.mask_of_ids([12, 25])] + [0] * 8
    n = 1 << 18
    full = gpu.api.equity_mc(holes, 0, 0, SEED, 1 << 33, n)        # counters above 2^32 exercise trial_hi
    parts = [gpu.api.equity_mc(holes, 0, 0, SEED, (1 << 33) + lo, cnt) for lo, cnt in [(0, 1), (1, 99_999), (100_000, n - 100_000)]]
    summed = tuple([sum(p[k][s] for p in parts) for s in range(9)] for k in range(3))
    assert summed == full
    rc, w, t, l = oracle.equity_mc(holes, 0, 0, SEED, 1 << 33, n)
    assert full == (w, t, l)


def test_equity_mc_other_shapes(gpu, oracle):
    rng = random.Random(5)
    api = gpu.api
    cases = [
        ([0, 0], 0, 0),                                               # everyone random
        ([api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50])], 0, 0),   # nothing random but the board
        ([api.mask_of_ids([0, 1]), 0, api.mask_of_ids([20, 33]), 0], api.mask_of_ids([5, 6, 7]), api.mask_of_ids([40, 41])),
        ([0] * 9, api.mask_of_ids([2, 16, 30, 44]), 0),
        ([api.mask_of_ids([3, 4])], 0, 0),
    ]
    for holes, board, dead in cases:
        seed = rng.getrandbits(64)
        got = api.equity_mc(holes, board, dead, seed, 17, 50_000)
        rc, w, t, l = oracle.equity_mc(holes, board, dead, seed, 17, 50_000)
        assert rc == 0 and got == (w, t, l)


def test_equity_mc_every_seat_shape(gpu, oracle):
    """every (random seats, fixed seats) split of 1..9 seats -- one K4 instantiation per number of random seats -- on
    preflop / flop / turn / river boards, counts equal to the oracle's"""
    rng = random.Random(11)
    api = gpu.api
    for n_seats in range(1, 10):
        for n_fixed in range(0, n_seats + 1):
            n_known = rng.choice([0, 0, 3, 4, 5])
            ids = rng.sample(range(52), 2 * n_fixed + n_known + 2)
            fixed_at = set(rng.sample(range(n_seats), n_fixed))
            holes, k = [], 0
            for s in range(n_seats):
                if s in fixed_at:
                    holes.append(api.mask_of_ids(ids[2 * k:2 * k + 2]))
                    k += 1
                else:
                    holes.append(0)
            board = api.mask_of_ids(ids[2 * n_fixed:2 * n_fixed + n_known])
            dead = api.mask_of_ids(ids[-2:]) if rng.random() < 0.5 else 0
            seed, first, n = rng.getrandbits(64), rng.getrandbits(40), 12_000
            got = api.equity_mc(holes, board, dead, seed, first, n)
            rc, w, t, l = oracle.equity_mc(holes, board, dead, seed, first, n)
            assert rc == 0 and got == (w, t, l), (n_seats, n_fixed, n_known)


def test_equity_mc_converges_to_exhaustive(gpu):
    """AA vs KK by Monte Carlo vs the exact count: within 4 sigma (binomial), the tolerance the spec allows"""
    api = gpu.api
    holes = [api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50])]
    n = 1 << 24
    w, t, l = api.equity_mc(holes, 0, 0, 999, 0, n)
    p = SG["aa_kk_no_shared"]["win_a"] / 1712304
    sigma = (p * (1 - p) / n) ** 0.5
    assert abs(w[0] / n - p) < 4 * sigma


def test_config4_acceptance_checks(gpu):
    """BASELINE config 4: 9 seats, seat 0 holds A-A, eight random seats.  (i) seat-0 equity of the bench seed
    within 4 sigma (binomial) of an independent-seed run; (ii) seats 1-8 are exchangeable, so their win / tie / loss counts
    agree with one another within 5 sigma of their binomial spread; (iii) every seat's counts add up to the trials;
    (iv) a trial has exactly one winner or at least two tied seats"""
    api = gpu.api
    holes = [api.mask_of_ids([12, 25])] + [0] * 8
    n = 1 << 26
    wa, ta, la = api.equity_mc(holes, 0, 0, SEED, 0, n)
    wb, tb, lb = api.equity_mc(holes, 0, 0, 0xC0FFEE1234, 0, n)
    for w, t, l in ((wa, ta, la), (wb, tb, lb)):
        assert all(w[s] + t[s] + l[s] == n for s in range(9))
        assert sum(w) <= n and sum(t) >= 2 * (n - sum(w)) and sum(t) <= 9 * (n - sum(w))
    eq = lambda w, t: (w[0] + t[0] / 2) / n
    pa, pb = eq(wa, ta), eq(wb, tb)
    sigma = (pa * (1 - pa) / n) ** 0.5
    assert abs(pa - pb) < 4 * sigma * 2 ** 0.5                                  # difference of two independent estimates
    assert 0.34 < pa < 0.37                                                    # A-A against eight random hands
    for counts in (wa[1:], la[1:], ta[1:]):
        mean = sum(counts) / 8
        p = mean / n
        sd = (n * p * (1 - p)) ** 0.5
        assert max(abs(c - mean) for c in counts) < 5 * sd, counts


# ---- showdown (gameplay.clj:118-133) ------------------------------------------------------------------
def test_showdown_groups(gpu, oracle):
    import ctypes as C
    rng = random.Random(3)
    for n_seats in (1, 2, 5, 9):
        for n_board in (3, 4, 5):
            ids = rng.sample(range(52), 2 * n_seats + n_board)
            holes = [gpu.api.mask_of_ids(ids[2 * s:2 * s + 2]) for s in range(n_seats)]
            board = gpu.api.mask_of_ids(ids[2 * n_seats:])
            keys, winners, group_of, n_groups = gpu.api.showdown(board, holes)
            want = [oracle.lib().orc_eval_mask(board | h) for h in holes]
            assert keys == want
            g = (C.c_int * n_seats)()
            ng = oracle.lib().orc_group_winners((C.c_uint32 * n_seats)(*want), n_seats, g)
            assert list(g) == group_of and ng == n_groups
            assert winners == sum(1 << s for s in range(n_seats) if want[s] == max(want))
    # a split pot: both play the board
    board = gpu.api.mask_of_ids([8, 9, 10, 11, 12])                   # hearts 10..A: royal flush on board
    keys, winners, group_of, n_groups = gpu.api.showdown(board, [gpu.api.mask_of_ids([13, 14]), gpu.api.mask_of_ids([26, 27])])
    assert winners == 0b11 and group_of == [0, 0] and n_groups == 1 and gpu.api.key_decode(keys[0])[0] == 8
    with pytest.raises(gpu.engine.McError) as e:                      # pre-flop showdown: SURVEY Q3
        gpu.api.showdown(0, [gpu.api.mask_of_ids([13, 14])])
    assert e.value.code == gpu.engine.MC_ERR_TOO_FEW


def test_end_game_ranking_mirror(gpu):
    Card = gpu.card.Card
    community = [Card("hearts", 2), Card("diamonds", 2), Card("spades", 9), Card("clubs", 11), Card("hearts", 5)]
    hands = {"a": [Card("spades", 2), Card("clubs", 9)], "b": [Card("hearts", 14), Card("diamonds", 13)],
             "c": [Card("spades", 14), Card("clubs", 13)]}
    values, groups = gpu.gameplay.end_game_ranking(community, hands)
    assert values["a"] == [6, [2, 2, 2, 9, 9], []]
    assert values["b"] == values["c"] == [1, [2, 2], [14, 13, 11]]
    assert groups == [{"a"}, {"b", "c"}]


# ---- K5: preflop matrix cells (BASELINE config 5) ------------------------------------------------------
def test_preflop_cells(gpu, oracle):
    import ctypes as C
    g = SG["cell_AA_KK"]
    cells = [(168, 154), (154, 168), (168, 168), (13 * 12 + 11, 13 * 0 + 5)]     # AA-KK, KK-AA, AA-AA, AKs-72o
    got = gpu.api.preflop_cells(cells)
    assert got[0].tolist() == [g["win"], g["tie"], g["loss"]]
    assert got[1].tolist() == [g["loss"], g["tie"], g["win"]]
    assert got[2][0] == got[2][2] and got[2].sum() == 6 * 1712304
    wtl = (C.c_int64 * 3)()
    assert oracle.lib().orc_preflop_cell(cells[3][0], cells[3][1], 0, wtl) == 0
    assert got[3].tolist() == list(wtl)


def test_preflop_matrix_full_properties(gpu):
    """BASELINE config 5: the whole 169 x 169 matrix by exhaustive boards; checked through the golden AA-KK cell,
    antisymmetry, row totals, and additivity of the per-rank parts (SURVEY 8d cfg 5)."""
    api = gpu.api
    m = api.preflop_matrix()
    aa, kk = api.hole_class(14, 14, False), api.hole_class(13, 13, False)
    assert m[aa, kk].tolist() == [50371344, 267444, 11004156]          # win, tie, loss of AA
    assert np.array_equal(m[:, :, 0], m[:, :, 2].T) and np.array_equal(m[:, :, 1], m[:, :, 1].T)
    boards = 1712304
    for cls in range(169):
        a, b = divmod(cls, 13)
        combos = 6 if a == b else (4 if a > b else 12)
        assert int(m[cls].sum()) == combos * 1225 * boards             # C(50,2) opposing holes x C(48,5) boards
    p0, n0 = api.preflop_matrix_part(0, 3)
    p1, n1 = api.preflop_matrix_part(1, 3)
    p2, n2 = api.preflop_matrix_part(2, 3)
    assert np.array_equal(p0 + p1 + p2, m) and n0 + n1 + n2 > 40000


def test_preflop_matrix_by_boards_equals_brute_force(gpu):
    """the per-board sort-and-count formulation (134,459 suit-canonical boards, 1,081 holes evaluated
    once per board) must reproduce the brute-force matrix of K5 cell by cell, also when dealt out in parts."""
    api = gpu.api
    want = api.preflop_matrix()
    got = api.preflop_matrix_by_boards()
    assert np.array_equal(got, want)
    w0, n0 = api.preflop_wins_by_boards(0, 2)
    w1, n1 = api.preflop_wins_by_boards(1, 2)
    assert n0 + n1 == 134459
    assert np.array_equal(api.preflop_matrix_from_wins(w0 + w1), want)


def test_abi_is_thread_safe(gpu, oracle):
    """the reference calls the evaluator from one go-block per room (board.clj:122-129): many host threads at once"""
    import threading
    api = gpu.api
    masks = oracle.random_hands(SEED + 3, 0, 20000)
    want = oracle.eval_batch(masks, 0)
    holes = [api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50])]
    errors = []

    def worker(k):
        try:
            for it in range(20):
                lo = (k * 997 + it * 131) % 15000
                if not np.array_equal(api.eval7_batch(masks[lo:lo + 4000]), want[lo:lo + 4000]):
                    errors.append("batch %d/%d" % (k, it))
                w, t, l, nb = api.equity_exhaustive(holes)
                if (w[0], w[1], t[0]) != (1388072, 318176, 6056):
                    errors.append("equity %d/%d" % (k, it))
                if api.key_decode(api.eval_cards([0, 0, 0, 0, 0], [9, 8, 7, 6, 5])) != [8, [9, 8, 7, 6, 5], []]:
                    errors.append("single %d/%d" % (k, it))
        except Exception as e:                      # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(6)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors[:5]


def test_end_game_one_call(gpu, oracle):
    """mc_end_game = end-game of gameplay.clj:122-133: ranking, tie groups and update-stacks (:104-116) in one call,
    against the oracle's showdown grouping + literal update-stacks on random tables and side-pot structures"""
    import ctypes as C
    rng = random.Random(17)
    for _ in range(60):
        n_seats, n_board = rng.randint(2, 9), rng.choice([3, 4, 5, 5, 5])
        ids = rng.sample(range(52), 2 * n_seats + n_board)
        if rng.random() < 0.3:                                          # force ties: everybody plays a made board
            board_ids, rest = [8, 9, 10, 11, 12], [i for i in range(13, 52)]
            ids = rng.sample(rest, 2 * n_seats) + board_ids
            n_board = 5
        holes = [gpu.api.mask_of_ids(ids[2 * s:2 * s + 2]) for s in range(n_seats)]
        board = gpu.api.mask_of_ids(ids[2 * n_seats:])
        pots = [(rng.randint(1, 500), rng.randint(1, n_seats), rng.sample(range(n_seats), rng.randint(1, n_seats)))
                for _ in range(rng.randint(0, 4))]
        keys, winners, group_of, n_groups, delta = gpu.api.end_game(board, holes, pots)
        want = [oracle.lib().orc_eval_mask(board | h) for h in holes]
        g = (C.c_int * n_seats)()
        ng = oracle.lib().orc_group_winners((C.c_uint32 * n_seats)(*want), n_seats, g)
        assert keys == want and list(g) == group_of and ng == n_groups
        sets = [sum(1 << s for s in range(n_seats) if g[s] == k) for k in range(ng)]
        stack = (C.c_longlong * 32)()
        rc = oracle.lib().orc_update_stacks((C.c_uint * ng)(*sets), ng, (C.c_longlong * max(1, len(pots)))(*[p[0] for p in pots]),
                                            (C.c_int * max(1, len(pots)))(*[p[1] for p in pots]),
                                            (C.c_uint * max(1, len(pots)))(*[sum(1 << s for s in p[2]) for p in pots]), len(pots), stack)
        assert rc == 0 and delta == list(stack[:n_seats])
        assert winners == sets[0]


def _random_pots(rng, n_seats, big=False):
    pots = []
    for _ in range(rng.randint(1, 8 if not big else 3)):
        players = rng.sample(range(n_seats), rng.randint(1, n_seats))
        if big:
            pots.append((rng.randint(1, 2 ** 26), rng.randint(1, 9), players))
        else:
            pots.append((rng.randint(0, 400), rng.randint(0, n_seats), players))
    return pots


def test_equity_ev_exhaustive_vs_oracle(gpu, oracle):
    """chips per seat under update-stacks (gameplay.clj:104-116) summed over every board, bit-exact against
    the oracle (group-winners + the literal update-stacks per board); counts equal to the plain call"""
    rng = random.Random(23)
    api = gpu.api
    for n_seats, n_board, n_dead, big in [(2, 0, 0, False), (2, 3, 0, True), (3, 3, 2, False), (9, 3, 0, False), (4, 4, 1, True),
                                          (6, 5, 0, False), (3, 2, 0, False), (5, 4, 0, False), (9, 4, 0, True)]:
        ids = rng.sample(range(52), 2 * n_seats + n_board + n_dead)
        holes = [api.mask_of_ids(ids[2 * s:2 * s + 2]) for s in range(n_seats)]
        board = api.mask_of_ids(ids[2 * n_seats:2 * n_seats + n_board])
        dead = api.mask_of_ids(ids[2 * n_seats + n_board:])
        pots = _random_pots(rng, n_seats, big)
        w, t, l, chips, nb = api.equity_exhaustive_ev(holes, pots, board, dead)
        rc, wo, to, lo, co, nbo = oracle.equity_exhaustive_ev(holes, pots, board, dead)
        assert rc == 0 and (w, t, l, chips, nb) == (wo, to, lo, co, nbo), (n_seats, n_board, pots)
        assert (w, t, l, nb) == api.equity_exhaustive(holes, board, dead)
    # AA vs KK, main pot 2 x 10 and a side pot only KK plays for: by hand from the golden counts
    holes = [api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50])]
    w, t, l, chips, nb = api.equity_exhaustive_ev(holes, [(10, 2, [0, 1]), (5, 1, [1])])
    assert chips == [20 * 1388072 + 10 * 6056, 20 * 318176 + 10 * 6056 + 5 * 1712304]
    E = gpu.engine
    for bad in ([], [(10, 2, [])], [(10, 2, [2])], [(2 ** 30, 2, [0, 1])], [(1, 1, [0])] * 9, [(-1, 2, [0])]):
        with pytest.raises(E.McError) as e:
            api.equity_exhaustive_ev(holes, bad)
        assert e.value.code == E.MC_ERR_INVALID


def test_equity_ev_monte_carlo_vs_oracle(gpu, oracle):
    """the same per Monte Carlo trial (K4's EV variant): every number of random seats, preflop to river, small and
    huge pots (the huge ones force the per-thread chip accumulators to fold every few trials), bit-exact vs the oracle;
    counts equal to the plain call's"""
    rng = random.Random(29)
    api = gpu.api
    for n_seats in range(1, 10):
        for n_fixed in sorted({0, 1, n_seats // 2, n_seats}):
            n_known = rng.choice([0, 0, 3, 4, 5])
            ids = rng.sample(range(52), 2 * n_fixed + n_known)
            fixed_at = set(rng.sample(range(n_seats), n_fixed))
            holes, k = [], 0
            for s in range(n_seats):
                if s in fixed_at:
                    holes.append(api.mask_of_ids(ids[2 * k:2 * k + 2]))
                    k += 1
                else:
                    holes.append(0)
            board = api.mask_of_ids(ids[2 * n_fixed:])
            pots = _random_pots(rng, n_seats, big=rng.random() < 0.3)
            seed, first, n = rng.getrandbits(64), rng.getrandbits(40), 20_000
            got = api.equity_mc_ev(holes, pots, board, 0, seed, first, n)
            rc, w, t, l, ch = oracle.equity_mc_ev(holes, pots, board, 0, seed, first, n)
            assert rc == 0 and got == (w, t, l, ch), (n_seats, n_fixed, n_known, pots)
            assert got[:3] == api.equity_mc(holes, board, 0, seed, first, n)
    # 9 seats, three nested pots, enough trials that every thread runs many: split invariance of the chip sums
    holes = [api.mask_of_ids([12, 25])] + [0] * 8
    pots = [(100, 9, list(range(9))), (250, 6, [0, 2, 3, 5, 6, 8]), (400, 3, [0, 3, 8])]
    full = api.equity_mc_ev(holes, pots, 0, 0, SEED, 0, 1 << 21)
    a = api.equity_mc_ev(holes, pots, 0, 0, SEED, 0, 700_001)
    b = api.equity_mc_ev(holes, pots, 0, 0, SEED, 700_001, (1 << 21) - 700_001)
    assert all([x + y for x, y in zip(a[k], b[k])] == full[k] for k in range(4))
    rc, w, t, l, ch = oracle.equity_mc_ev(holes, pots, 0, 0, SEED, 0, 1 << 18)
    assert api.equity_mc_ev(holes, pots, 0, 0, SEED, 0, 1 << 18) == (w, t, l, ch)


def test_equity_mc_ranges_vs_oracle(gpu, oracle):
    """SURVEY 8f.3 tail: Monte Carlo with a weighted hole range per seat (deal spec R1) -- bit-exact against the oracle's
    restatement of the spec for 1..9 seats, mixed fixed / ranged seats, preflop to river, narrow ranges that collide
    often, dead cards; split invariance; heads-up within 4 sigma of the exhaustive range-vs-range call; void trials"""
    api = gpu.api
    rng = random.Random(41)
    holes = [(a, b) for b in range(52) for a in range(b)]
    for n_seats in (1, 2, 3, 6, 9):
        for case in range(3):
            n_known = rng.choice([0, 3, 4, 5])
            n_fixed = rng.randint(0, n_seats - 1) if case else 0
            ids = rng.sample(range(52), 2 * n_fixed + n_known + 2)
            fixed_at = set(rng.sample(range(n_seats), n_fixed))
            hm, weights, k = [], [], 0
            for s in range(n_seats):
                if s in fixed_at:
                    hm.append(api.mask_of_ids(ids[2 * k:2 * k + 2]))
                    weights.append(None)
                    k += 1
                else:
                    hm.append(0)
                    w = [0] * 1326
                    for a, b in rng.sample(holes, rng.choice([3, 20, 200, 1326])):
                        w[api.hole_index(a, b)] = rng.randint(1, 32767 if case == 2 else 9)
                    weights.append(w)
            board = api.mask_of_ids(ids[2 * n_fixed:2 * n_fixed + n_known])
            dead = api.mask_of_ids(ids[-2:]) if rng.random() < 0.5 else 0
            seed, first, n = rng.getrandbits(64), rng.getrandbits(40), 20_000
            rc, w, t, l, void = oracle.equity_mc_ranges(weights, hm, board, dead, seed, first, n)
            if rc != 0:                                             # a range emptied by the known cards
                with pytest.raises(gpu.engine.McError):
                    api.equity_mc_ranges(weights, hm, board, dead, seed, first, n)
                continue
            got = api.equity_mc_ranges(weights, hm, board, dead, seed, first, n)
            assert got == (w, t, l, void), (n_seats, case, n_known, n_fixed)
            assert all(a + b + c + void == n for a, b, c in zip(w, t, l))
    # heads-up: Monte Carlo against the exact weighted equity of mc_equity_ranges, split invariance
    wa, wb = [0] * 1326, [0] * 1326
    for a, b in rng.sample(holes, 40):
        wa[api.hole_index(a, b)] = rng.randint(1, 9)
    for a, b in rng.sample(holes, 40):
        wb[api.hole_index(a, b)] = rng.randint(1, 9)
    board = api.mask_of_ids([3, 17, 44])
    exact, total = api.equity_ranges(wa, wb, board, 0)
    n = 1 << 22
    w, t, l, void = api.equity_mc_ranges([wa, wb], [0, 0], board, 0, 99, 0, n)
    assert void == 0 and w[0] == l[1] and t[0] == t[1]
    p = exact[0] / total
    assert abs(w[0] / n - p) < 4 * (p * (1 - p) / n) ** 0.5
    a = api.equity_mc_ranges([wa, wb], [0, 0], board, 0, 99, 0, 1_000_001)
    b = api.equity_mc_ranges([wa, wb], [0, 0], board, 0, 99, 1_000_001, n - 1_000_001)
    assert [x + y for x, y in zip(a[0], b[0])] == w and a[3] + b[3] == void
    # two ranges that can never be dealt together: every trial is void
    only, other = [0] * 1326, [0] * 1326
    only[api.hole_index(12, 25)] = 3
    other[api.hole_index(12, 38)] = 1
    assert api.equity_mc_ranges([only, other], [0, 0], 0, 0, 1, 0, 300) == ([0, 0], [0, 0], [0, 0], 300)


def test_lifecycle_is_safe_against_calls_in_flight(gpu, oracle):
    """the header's promise: mc_init* / mc_shutdown wait for calls in flight and calls never run on a context that is
    being torn down.  Four threads hammer compute calls (batch, single-hand ticket path, equity) while a fifth cycles
    mc_shutdown / mc_init; every call either returns the right answer or fails with MC_ERR_NOT_INIT."""
    import threading
    import time
    E, api = gpu.engine, gpu.api
    L = E.lib()
    masks = oracle.random_hands(SEED + 5, 0, 30000)
    want = oracle.eval_batch(masks, 0)
    holes = [api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50])]
    errors, done, stop = [], [0], threading.Event()

    def call(fn):
        try:
            return fn()
        except E.McError as e:
            if e.code != E.MC_ERR_NOT_INIT:
                errors.append(repr(e))
            return None

    def worker(k):
        it = 0
        while not stop.is_set():
            it += 1
            lo = (k * 997 + it * 131) % 20000
            got = call(lambda: api.eval7_batch(masks[lo:lo + 5000]))
            if got is not None and not np.array_equal(got, want[lo:lo + 5000]):
                errors.append("batch %d/%d" % (k, it))
            got = call(lambda: api.eval_cards([0, 0, 0, 0, 0], [9, 8, 7, 6, 5]))
            if got is not None and got != 0x898765:
                errors.append("single %d/%d: %x" % (k, it, got))
            got = call(lambda: api.equity_exhaustive(holes))
            if got is not None and (got[0][0], got[0][1], got[1][0]) != (1388072, 318176, 6056):
                errors.append("equity %d/%d" % (k, it))
            got = call(lambda: api.showdown(api.mask_of_ids([8, 9, 10, 11, 12]), [api.mask_of_ids([13, 14]), api.mask_of_ids([26, 27])]))
            if got is not None and (got[1], got[2]) != (3, [0, 0]):
                errors.append("showdown %d/%d" % (k, it))
            done[0] += 1

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    for th in threads:
        th.start()
    try:
        for _ in range(6):
            time.sleep(0.05)
            L.mc_shutdown()
            time.sleep(0.02)
            E.check(L.mc_init(0, 0))
    finally:
        time.sleep(0.1)
        stop.set()
        for th in threads:
            th.join()
        E.check(L.mc_init(0, 0))                                      # leave the session fixture's engine initialised
    assert not errors, errors[:5]
    assert done[0] > 0


def test_multi_device_split_equals_one_device(gpu, oracle):
    """the host-buffer calls split their work over every device of the library instance (single-process mode for a JVM
    host): mc_init(0) on a multi-GPU box must give exactly what mc_init_devices({0}) gives.  Skipped on a 1-GPU box."""
    import ctypes as C
    import torch
    E, api = gpu.engine, gpu.api
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least two devices")
    masks = oracle.random_hands(SEED + 31, 0, 3_000_001)
    holes9 = [api.mask_of_ids([12, 25])] + [0] * 8

    def everything():
        return (api.eval7_batch(masks), api.eval_batch_colex32(api.colex_rank(masks)), api.eval_batch_packed48(api.pack48(masks)),
                api.enumerate7(5, 3_000_000), api.equity_exhaustive([api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50])]),
                api.equity_mc(holes9, 0, 0, SEED, 7, 1 << 20), api.preflop_cells([(168, 154), (100, 7)]).tolist())
    try:
        assert E.init(devices=[0]) == 1
        one = everything()
        assert E.init() == torch.cuda.device_count() and E.lib().mc_collective() == 0
        many = everything()
        # the same with the counts summed by one ncclAllReduce per device inside the library (MC_INIT_NCCL)
        assert E.init(flags=E.INIT_NCCL) == torch.cuda.device_count() and E.lib().mc_collective() == 1
        nccl = everything()
    finally:
        E.init()
    for other in (many, nccl):
        for a, b in zip(one, other):
            if isinstance(a, np.ndarray):
                assert np.array_equal(a, b)
            else:
                assert a == b


def test_equity_ranges_vs_oracle(gpu, oracle):
    """range vs range on flop / turn / river / preflop boards == sum of the oracle's exhaustive
    equities over the weighted disjoint hole pairs"""
    import random
    api = gpu.api
    rnd = random.Random(3)
    for n_board, n_a, n_b in ((3, 25, 30), (4, 40, 40), (5, 60, 50), (0, 3, 3), (3, 1326, 8)):
        cards = rnd.sample(range(52), n_board + 2)
        board, dead = api.mask_of_ids(cards[:n_board]), api.mask_of_ids(cards[n_board:])
        holes = [(a, b) for b in range(52) for a in range(b)]
        wa, wb = [0] * 1326, [0] * 1326
        for a, b in rnd.sample(holes, n_a):
            wa[api.hole_index(a, b)] = rnd.randint(1, 9)
        for a, b in rnd.sample(holes, n_b):
            wb[api.hole_index(a, b)] = rnd.randint(1, 9)
        got, total = api.equity_ranges(wa, wb, board, dead)
        want, wtot = [0, 0, 0], 0
        la = [(wa[api.hole_index(a, b)], api.mask_of_ids([a, b])) for a, b in holes if wa[api.hole_index(a, b)]]
        lb = [(wb[api.hole_index(a, b)], api.mask_of_ids([a, b])) for a, b in holes if wb[api.hole_index(a, b)]]
        for fa, ha in la:
            if ha & (board | dead):
                continue
            for fb, hb in lb:
                if hb & (board | dead | ha):
                    continue
                rc, w, t, l, nb = oracle.equity_exhaustive([ha, hb], board, dead)
                want[0] += fa * fb * w[0]; want[1] += fa * fb * t[0]; want[2] += fa * fb * l[0]; wtot += fa * fb * nb
        assert got == want and total == wtot and sum(got) == total
    # AA vs KK as two one-class ranges, preflop = the golden class cell
    wa, wb = [0] * 1326, [0] * 1326
    for s in range(4):
        for t in range(s):
            wa[api.hole_index(13 * s + 12, 13 * t + 12)] = 1
            wb[api.hole_index(13 * s + 11, 13 * t + 11)] = 1
    got, total = api.equity_ranges(wa, wb)
    assert got == [50371344, 267444, 11004156] and total == 36 * 1712304
