"""numpy-facing wrappers over the C ABI (host buffers in, host buffers out).  These are what the Python
mirror of the reference's namespaces (hand_evaluator.py, gameplay.py), the tests and bench.py's end-to-end
leg call; each is a thin argument-marshalling layer over one mc_* entry point."""
import ctypes as C

import numpy as np

from . import engine as E

HIT_LEN = (5, 2, 4, 3, 5, 5, 5, 4, 5)        # per category 0..8 (hand_evaluator.clj:100-133)


def card_bit(card_id):
    return int(E.lib().mc_card_bit(int(card_id)))


def mask_of_ids(ids):
    m = 0
    for i in ids:
        b = card_bit(i)
        if not b:
            raise ValueError("card id out of range: %r" % (i,))
        m |= b
    return m


def key_decode(key):
    """nibble key -> [cat, [hit...], [kick...]] exactly as the Clojure value (cat 6 slot 1 as a vector)."""
    cat, nh, nk = C.c_int32(), C.c_int32(), C.c_int32()
    hit, kick = (C.c_int32 * 5)(), (C.c_int32 * 5)()
    E.check(E.lib().mc_key_decode(int(key), C.byref(cat), hit, C.byref(nh), kick, C.byref(nk)))
    return [cat.value, list(hit[:nh.value]), list(kick[:nk.value])]


def key_packed_to_nibble(key):
    return int(E.lib().mc_key_packed_to_nibble(int(key)))


def eval_cards(suit_ids, ranks):
    E.ensure_init()
    n = len(ranks)
    s = (C.c_int32 * max(n, 1))(*suit_ids)
    r = (C.c_int32 * max(n, 1))(*ranks)
    key = C.c_uint32(0)
    E.check(E.lib().mc_eval_cards(s, r, n, C.byref(key)))
    return key.value


def eval7_batch(masks, key_format=E.KEY_NIBBLE, out=None):
    E.ensure_init()
    masks = np.ascontiguousarray(masks, dtype=np.uint64)
    keys = out if out is not None else np.empty(masks.shape[0], dtype=np.uint32)
    E.check(E.lib().mc_eval7_batch_fmt(masks.ctypes.data, masks.shape[0], keys.ctypes.data, key_format))
    return keys


def pack48(masks):
    """uint64 card masks (<= 7 cards each) -> the 6-byte wire format of mc_eval_batch_packed48 (host bit shuffling
    in numpy: seven 6-bit card positions per hand, lowest card first, unused fields 0)"""
    m = np.ascontiguousarray(masks, dtype=np.uint64).copy()
    h = np.zeros(m.shape[0], dtype=np.uint64)
    one = np.uint64(1)
    for k in range(7):
        low = m & (~m + one)                                   # lowest set bit (0 when the hand is used up)
        pos = np.zeros(m.shape[0], dtype=np.uint64)
        for b in (32, 16, 8, 4, 2, 1):                         # log2 of a power of two, branch-free
            big = (low >> np.uint64(b)) != 0
            pos += np.where(big, np.uint64(b), np.uint64(0))
            low = np.where(big, low >> np.uint64(b), low)
        h |= pos << np.uint64(6 * k)
        m &= m - one
    if np.any(m):
        raise ValueError("a hand holds more than 7 cards")
    return np.ascontiguousarray(h.view(np.uint8).reshape(-1, 8)[:, :6]).reshape(-1)


def eval_batch_packed48(packed, out=None):
    """packed: uint8 array of 6*n bytes (pack48) -> nibble keys"""
    E.ensure_init()
    packed = np.ascontiguousarray(packed, dtype=np.uint8)
    if packed.size % 6:
        raise ValueError("packed48 hands are 6 bytes each")
    n = packed.size // 6
    keys = out if out is not None else np.empty(n, dtype=np.uint32)
    E.check(E.lib().mc_eval_batch_packed48(packed.ctypes.data, n, keys.ctypes.data))
    return keys


def colex_rank(masks):
    """uint64 masks of 7-card hands of the deck -> uint32 colex ranks over card ids (sum_i C(c_i, i+1)); host
    arithmetic in numpy: what a caller of mc_eval_batch_colex32 computes from its cards"""
    from math import comb
    m = np.ascontiguousarray(masks, dtype=np.uint64).copy()
    binom = np.array([[comb(x, k) for x in range(52)] for k in range(8)], dtype=np.uint64)
    rank = np.zeros(m.shape[0], dtype=np.uint64)
    one = np.uint64(1)
    for k in range(1, 8):
        low = m & (~m + one)
        pos = np.zeros(m.shape[0], dtype=np.int64)
        for b in (32, 16, 8, 4, 2, 1):
            big = (low >> np.uint64(b)) != 0
            pos += np.where(big, b, 0)
            low = np.where(big, low >> np.uint64(b), low)
        if np.any(m == 0):
            raise ValueError("a hand holds fewer than 7 cards")
        cid = 13 * (pos >> 4) + (pos & 15) - 2
        if np.any((cid < 0) | (cid > 51) | ((pos & 15) < 2) | ((pos & 15) > 14)):
            raise ValueError("a bit that is not one of the 52 cards")
        rank += binom[k][cid]
        m &= m - one
    if np.any(m):
        raise ValueError("a hand holds more than 7 cards")
    return rank.astype(np.uint32)


def pack27(ranks):
    """uint32 colex ranks (< 2^27) -> the 27-bit little-endian bit stream of mc_eval_batch_colex27 (uint32 words,
    ceil(27 n / 32) of them): hand j occupies bits [27 j, 27 j + 27)"""
    r = np.ascontiguousarray(ranks, dtype=np.uint32)
    if np.any(r >> np.uint32(27)):
        raise ValueError("a rank does not fit 27 bits")
    n = r.shape[0]
    groups = (n + 31) // 32
    pad = np.zeros(groups * 32, dtype=np.uint64)
    pad[:n] = r
    pad = pad.reshape(groups, 32)
    out = np.zeros((groups, 28), dtype=np.uint64)                      # (word 27 only catches the shifted-out zeros)
    for j in range(32):
        w, sh = (27 * j) >> 5, (27 * j) & 31
        v = pad[:, j] << np.uint64(sh)
        out[:, w] |= v & np.uint64(0xFFFFFFFF)
        out[:, w + 1] |= v >> np.uint64(32)
    words = out[:, :27].astype(np.uint32).reshape(-1)
    return np.ascontiguousarray(words[:(27 * n + 31) // 32])


def eval_batch_colex27(stream, n, out=None):
    """the 27-bit stream of n hands (pack27) -> uint16 class ids"""
    E.ensure_init()
    stream = np.ascontiguousarray(stream, dtype=np.uint32)
    if stream.shape[0] < (27 * n + 31) // 32:
        raise ValueError("the stream is shorter than ceil(27 n / 32) words")
    cls = out if out is not None else np.empty(n, dtype=np.uint16)
    E.check(E.lib().mc_eval_batch_colex27(stream.ctypes.data, n, cls.ctypes.data))
    return cls


def eval_batch_colex32(ranks, out=None):
    """uint32 colex ranks of 7-card hands -> uint16 class ids (class_keys()[id] = nibble key)"""
    E.ensure_init()
    ranks = np.ascontiguousarray(ranks, dtype=np.uint32)
    cls = out if out is not None else np.empty(ranks.shape[0], dtype=np.uint16)
    E.check(E.lib().mc_eval_batch_colex32(ranks.ctypes.data, ranks.shape[0], cls.ctypes.data))
    return cls


def eval_batch(hands, in_format, key_format=E.KEY_NIBBLE, stage_cards=0):
    """any wire format in (IN_MASK64 uint64, IN_PACKED48 6 bytes, IN_COLEX32 uint32), keys (uint32) or class ids
    (uint16, KEY_CLASS16) out"""
    E.ensure_init()
    fmt = {E.IN_MASK64: (np.uint64, 1), E.IN_PACKED48: (np.uint8, 6), E.IN_COLEX32: (np.uint32, 1)}.get(in_format)      # (colex27: eval_batch_colex27)
    hands = np.ascontiguousarray(hands, dtype=fmt[0]) if fmt else np.ascontiguousarray(hands)   # (an unknown format is the library's to refuse)
    n = hands.size // (fmt[1] if fmt else 1)
    out = np.empty(n, dtype=np.uint16 if key_format == E.KEY_CLASS16 else np.uint32)
    E.check(E.lib().mc_eval_batch(hands.ctypes.data, in_format, n, out.ctypes.data, key_format, stage_cards))
    return out


def class_keys():
    """class id -> nibble key (7,462 entries, ascending)"""
    E.ensure_init()
    out = np.empty(E.N_HAND_CLASSES, dtype=np.uint32)
    E.check(E.lib().mc_class_keys(out.ctypes.data, E.N_HAND_CLASSES))
    return out


def random_hands(seed, first, n, n_cards=7):
    E.ensure_init()
    out = np.empty(n, dtype=np.uint64)
    E.check(E.lib().mc_random_hands(seed, first, n, n_cards, out.ctypes.data))
    return out


def showdown(board_mask, hole_masks):
    """-> (keys per seat, winner bitmask, tie-group index per seat (0 = best), number of groups)"""
    E.ensure_init()
    n = len(hole_masks)
    h = (C.c_uint64 * max(n, 1))(*hole_masks)
    keys, groups, ng, win = (C.c_uint32 * 9)(), (C.c_int32 * 9)(), C.c_int32(0), C.c_uint32(0)
    E.check(E.lib().mc_showdown_groups(board_mask, h, n, keys, groups, C.byref(ng), C.byref(win)))
    return list(keys[:n]), win.value, list(groups[:n]), ng.value


def _pot_arrays(pots):
    k = len(pots)
    bet = (C.c_int64 * max(k, 1))(*[p[0] for p in pots])
    cnt = (C.c_int32 * max(k, 1))(*[p[1] for p in pots])
    msk = (C.c_uint32 * max(k, 1))(*[sum(1 << s for s in p[2]) for p in pots])
    return bet, cnt, msk, k


def end_game(board_mask, hole_masks, pots):
    """end-game in one call (gameplay.clj:122-133): pots = [(bet, n, [seats])] ->
    (keys, winner bitmask, tie group per seat, number of groups, chips won per seat)"""
    E.ensure_init()
    n = len(hole_masks)
    h = (C.c_uint64 * max(n, 1))(*hole_masks)
    keys, groups, ng, win = (C.c_uint32 * 9)(), (C.c_int32 * 9)(), C.c_int32(0), C.c_uint32(0)
    bet, cnt, msk, k = _pot_arrays(pots)
    delta = (C.c_int64 * 9)()
    E.check(E.lib().mc_end_game(board_mask, h, n, bet, cnt, msk, k, keys, groups, C.byref(ng), C.byref(win), delta))
    return list(keys[:n]), win.value, list(groups[:n]), ng.value, list(delta[:n])


def settle_pots(group_of, pots):
    """update-stacks (gameplay.clj:104-116): group_of[seat] = tie group (0 best, -1 = not at the showdown);
    pots = [(bet, n, [seats])] -> chips won per seat"""
    n = len(group_of)
    g = (C.c_int32 * n)(*group_of)
    bet, cnt, msk, k = _pot_arrays(pots)
    delta = (C.c_int64 * n)()
    E.check(E.lib().mc_settle_pots(g, n, bet, cnt, msk, k, delta))
    return list(delta)


def enumerate7(first=None, count=None, key_sum=True):
    E.ensure_init()
    hist, ks = (C.c_uint64 * 9)(), C.c_uint64(0)
    ksp = C.byref(ks) if key_sum else None
    if first is None:
        E.check(E.lib().mc_enumerate7(hist, ksp))
    else:
        E.check(E.lib().mc_enumerate7_range(first, count, hist, ksp))
    return list(hist), (ks.value if key_sum else None)


def _seat_arrays(n):
    return (C.c_int64 * 9)(), (C.c_int64 * 9)(), (C.c_int64 * 9)()


def equity_exhaustive(hole_masks, board_mask=0, dead_mask=0, first=None, count=None):
    E.ensure_init()
    n = len(hole_masks)
    h = (C.c_uint64 * max(n, 1))(*hole_masks)
    w, t, l = _seat_arrays(n)
    nb = C.c_uint64(0)
    if first is None:
        E.check(E.lib().mc_equity_exhaustive(h, n, board_mask, dead_mask, w, t, l, C.byref(nb)))
    else:
        E.check(E.lib().mc_equity_exhaustive_range(h, n, board_mask, dead_mask, first, count, w, t, l))
        nb.value = count
    return list(w[:n]), list(t[:n]), list(l[:n]), nb.value


def equity_mc(hole_masks, board_mask, dead_mask, seed, first_trial, n_trials):
    E.ensure_init()
    n = len(hole_masks)
    h = (C.c_uint64 * max(n, 1))(*hole_masks)
    w, t, l = _seat_arrays(n)
    E.check(E.lib().mc_equity_mc(h, n, board_mask, dead_mask, seed, first_trial, n_trials, w, t, l))
    return list(w[:n]), list(t[:n]), list(l[:n])


def equity_exhaustive_ev(hole_masks, pots, board_mask=0, dead_mask=0):
    """exhaustive equity + the reference's pot settlement per board: pots = [(bet, n, [seats])] ->
    (win, tie, loss, chips per seat summed over all boards, number of boards)"""
    E.ensure_init()
    n = len(hole_masks)
    h = (C.c_uint64 * max(n, 1))(*hole_masks)
    w, t, l = _seat_arrays(n)
    chips, nb = (C.c_int64 * 9)(), C.c_uint64(0)
    bet, cnt, msk, k = _pot_arrays(pots)
    E.check(E.lib().mc_equity_exhaustive_ev(h, n, board_mask, dead_mask, bet, cnt, msk, k, w, t, l, chips, C.byref(nb)))
    return list(w[:n]), list(t[:n]), list(l[:n]), list(chips[:n]), nb.value


def equity_mc_ev(hole_masks, pots, board_mask, dead_mask, seed, first_trial, n_trials):
    """Monte Carlo equity + the reference's pot settlement per trial -> (win, tie, loss, chips per seat summed over the trials)"""
    E.ensure_init()
    n = len(hole_masks)
    h = (C.c_uint64 * max(n, 1))(*hole_masks)
    w, t, l = _seat_arrays(n)
    chips = (C.c_int64 * 9)()
    bet, cnt, msk, k = _pot_arrays(pots)
    E.check(E.lib().mc_equity_mc_ev(h, n, board_mask, dead_mask, seed, first_trial, n_trials, bet, cnt, msk, k, w, t, l, chips))
    return list(w[:n]), list(t[:n]), list(l[:n]), list(chips[:n])


def hole_index(card_a, card_b):
    """colex rank of a two-card hole over card ids (0..1325)"""
    c0, c1 = min(card_a, card_b), max(card_a, card_b)
    return c0 + c1 * (c1 - 1) // 2


def equity_ranges(weight_a, weight_b, board_mask=0, dead_mask=0):
    """heads-up range vs range: weights per hole index (1326 each) -> ([win, tie, loss] of A, total), weighted"""
    E.ensure_init()
    wa = (C.c_int32 * 1326)(*weight_a)
    wb = (C.c_int32 * 1326)(*weight_b)
    out, tot = (C.c_int64 * 3)(), C.c_int64(0)
    E.check(E.lib().mc_equity_ranges(wa, wb, board_mask, dead_mask, out, C.byref(tot)))
    return list(out), tot.value


def equity_mc_ranges(weights, hole_masks, board_mask, dead_mask, seed, first_trial, n_trials):
    """Monte Carlo with a hole range per seat: weights = one list of 1326 ints per seat (None / ignored for a fixed seat),
    hole_masks[s] != 0 fixes seat s -> (win, tie, loss, void trials)"""
    E.ensure_init()
    n = len(hole_masks)
    w = np.zeros((max(n, 1), 1326), dtype=np.int32)
    for s in range(n):
        if weights[s] is not None:
            w[s] = np.asarray(weights[s], dtype=np.int32)
    h = (C.c_uint64 * max(n, 1))(*hole_masks)
    wi, ti, lo = _seat_arrays(n)
    void = C.c_uint64(0)
    E.check(E.lib().mc_equity_mc_ranges(w.ctypes.data, h, n, board_mask, dead_mask, seed, first_trial, n_trials, wi, ti, lo, C.byref(void)))
    return list(wi[:n]), list(ti[:n]), list(lo[:n]), void.value


def preflop_cells(cells):
    E.ensure_init()
    n = len(cells)
    a = (C.c_int32 * n)(*[c[0] for c in cells])
    b = (C.c_int32 * n)(*[c[1] for c in cells])
    out = np.zeros((n, 3), dtype=np.int64)
    E.check(E.lib().mc_preflop_cells(a, b, n, out.ctypes.data))
    return out


def preflop_matrix():
    E.ensure_init()
    out = np.zeros((E.N_CLASSES, E.N_CLASSES, 3), dtype=np.int64)
    E.check(E.lib().mc_preflop_matrix(out.ctypes.data))
    return out


def preflop_matrix_part(part, n_parts):
    """-> (this part's contribution to the 169x169x3 matrix, number of suit-canonical matchups it covered)"""
    E.ensure_init()
    out = np.zeros((E.N_CLASSES, E.N_CLASSES, 3), dtype=np.int64)
    n = C.c_int32(0)
    E.check(E.lib().mc_preflop_matrix_part(part, n_parts, out.ctypes.data, C.byref(n)))
    return out, n.value


def preflop_boards_count():
    """-> (number of suit-canonical 5-card boards, sum of their orbit sizes); host only, needs no GPU"""
    n, w = C.c_int32(0), C.c_int64(0)
    E.check(E.lib().mc_preflop_boards_count(C.byref(n), C.byref(w)))
    return n.value, w.value


def preflop_wins_by_boards(part=0, n_parts=1):
    """-> (this part's share of the 169x169 win matrix computed board by board, number of canonical boards it covered)"""
    E.ensure_init()
    out = np.zeros((E.N_CLASSES, E.N_CLASSES), dtype=np.int64)
    n = C.c_int32(0)
    E.check(E.lib().mc_preflop_wins_by_boards(part, n_parts, out.ctypes.data, C.byref(n)))
    return out, n.value


def preflop_matrix_from_wins(win):
    """full win matrix -> 169x169x3 {win, tie, loss} (host arithmetic)"""
    win = np.ascontiguousarray(win, dtype=np.int64)
    out = np.zeros((E.N_CLASSES, E.N_CLASSES, 3), dtype=np.int64)
    E.check(E.lib().mc_preflop_matrix_from_wins(win.ctypes.data, out.ctypes.data))
    return out


def preflop_matrix_by_boards():
    """the 169x169x3 matrix by the per-board sort-and-count formulation; equal to preflop_matrix() cell by cell"""
    E.ensure_init()
    out = np.zeros((E.N_CLASSES, E.N_CLASSES, 3), dtype=np.int64)
    E.check(E.lib().mc_preflop_matrix_by_boards(out.ctypes.data))
    return out


def hole_class(rank_a, rank_b, suited):
    """169-class id: x = rank-2; pair 13x+x; suited 13*hi+lo; offsuit 13*lo+hi"""
    hi, lo = max(rank_a, rank_b) - 2, min(rank_a, rank_b) - 2
    if hi == lo:
        return 13 * hi + hi
    return 13 * hi + lo if suited else 13 * lo + hi
