"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by the product package."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_DIR = os.path.join(os.path.dirname(_HERE), "oracle")
_LIB = None


class Card(C.Structure):
    _fields_ = [("suit", C.c_int), ("rank", C.c_int)]


class Value(C.Structure):
    _fields_ = [("cat", C.c_int), ("nhit", C.c_int), ("hit", C.c_int * 7),
                ("nkick", C.c_int), ("kick", C.c_int * 7), ("hit_is_lazyseq", C.c_int)]

    def as_list(self):
        return [self.cat, list(self.hit[:self.nhit]), list(self.kick[:self.nkick])]


def build(force=False):
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("ref_literal.c", "ref_fast.c", "mc_oracle.h")]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    L = C.CDLL(build())
    u64, u32, i64, i32 = C.c_uint64, C.c_uint32, C.c_int64, C.c_int
    P = C.POINTER
    L.orc_constant_coll.argtypes = [P(i32), i32]
    L.orc_straight_p.argtypes = [P(Card), i32]
    L.orc_flush_p.argtypes = [P(Card), i32]
    L.orc_n_of_a_kind_p.argtypes = [P(Card), i32, i32, P(Card)]
    L.orc_full_house_p.argtypes = [P(Card), i32]
    L.orc_two_pair_p.argtypes = [P(Card), i32]
    L.orc_evaluator.argtypes = [P(Card), P(Value)]
    L.orc_evaluator.restype = None
    L.orc_compare.argtypes = [P(Value), P(Value), P(i32)]
    L.orc_player_hand_value.argtypes = [P(Card), i32, P(Value), P(i32)]
    L.orc_group_winners.argtypes = [P(u32), i32, P(i32)]
    L.orc_update_stacks.argtypes = [P(C.c_uint), i32, P(C.c_longlong), P(C.c_int), P(C.c_uint), i32, P(C.c_longlong)]
    L.orc_value_to_key.argtypes = [P(Value)]
    L.orc_value_to_key.restype = u32
    L.orc_key_to_value.argtypes = [u32, P(Value)]
    L.orc_key_to_value.restype = None
    L.orc_rc5_key.argtypes = [P(i32), P(i32)]
    L.orc_rc5_key.restype = u32
    for f in (L.orc_best_of_subsets_key, L.orc_eval_mask):
        f.argtypes = [u64]
        f.restype = u32
    L.orc_eval_mask_literal.argtypes = [u64, P(i32)]
    L.orc_eval_mask_literal.restype = u32
    L.orc_eval_batch.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, i32, i32]
    L.orc_eval_batch.restype = None
    L.orc_enumerate.argtypes = [i32, i32, i32, u64, u64, P(u64), P(u64), P(u64)]
    L.orc_enumerate_crosscheck.argtypes = [i32, i32]
    L.orc_enumerate_crosscheck.restype = u64
    L.orc_equity_exhaustive.argtypes = [P(u64), i32, u64, u64, i32, P(i64), P(i64), P(i64), P(u64)]
    L.orc_philox4x32_10.argtypes = [P(u32), P(u32), P(u32)]
    L.orc_philox4x32_10.restype = None
    L.orc_random_hand_mask.argtypes = [u64, u64, i32]
    L.orc_random_hand_mask.restype = u64
    L.orc_random_hands.argtypes = [u64, u64, C.c_size_t, i32, C.c_void_p]
    L.orc_random_hands.restype = None
    L.orc_deal_trial.argtypes = [P(u64), i32, u64, u64, u64, u64, P(u64), P(u64)]
    L.orc_equity_mc.argtypes = [P(u64), i32, u64, u64, u64, u64, u64, i32, P(i64), P(i64), P(i64)]
    L.orc_preflop_cell.argtypes = [i32, i32, i32, P(i64)]
    _LIB = L
    return L


# ---- small pythonic helpers ---------------------------------------------------------------------
def cards(pairs):
    """[(suit, rank), ...] -> ctypes array of orc_card"""
    arr = (Card * len(pairs))()
    for i, (s, r) in enumerate(pairs):
        arr[i].suit, arr[i].rank = s, r
    return arr


def evaluator(pairs):
    v = Value()
    lib().orc_evaluator(cards(pairs), C.byref(v))
    return v


def player_hand_value(pairs):
    v, thr = Value(), C.c_int(0)
    rc = lib().orc_player_hand_value(cards(pairs), len(pairs), C.byref(v), C.byref(thr))
    return rc, v, thr.value


def mask_of(pairs):
    m = 0
    for s, r in pairs:
        m |= 1 << (16 * s + r)
    return m


def card_bit(card_id):
    return 1 << (16 * (card_id // 13) + card_id % 13 + 2)


def eval_batch(masks, which=0, threads=0):
    import numpy as np
    masks = np.ascontiguousarray(masks, dtype=np.uint64)
    keys = np.empty(masks.shape[0], dtype=np.uint32)
    lib().orc_eval_batch(masks.ctypes.data, masks.shape[0], keys.ctypes.data, which, threads)
    return keys


def random_hands(seed, first, n, n_cards=7):
    import numpy as np
    out = np.empty(n, dtype=np.uint64)
    lib().orc_random_hands(seed, first, n, n_cards, out.ctypes.data)
    return out


def enumerate_hands(k, which=0, threads=0, first=0, count=0, distinct=False):
    hist = (C.c_uint64 * 9)()
    ks, nd = C.c_uint64(0), C.c_uint64(0)
    rc = lib().orc_enumerate(k, which, threads, first, count, hist, C.byref(ks), C.byref(nd) if distinct else None)
    assert rc == 0
    return list(hist), ks.value, (nd.value if distinct else None)


def enumerate_masks(k, first, count, threads=0):
    import numpy as np
    out = np.empty(count, dtype=np.uint64)
    L = lib()
    L.orc_enumerate_masks.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p]
    assert L.orc_enumerate_masks(k, first, count, threads, out.ctypes.data) == 0
    return out


def equity_exhaustive(holes, board=0, dead=0, threads=0):
    n = len(holes)
    h = (C.c_uint64 * n)(*holes)
    w, t, l = (C.c_int64 * n)(), (C.c_int64 * n)(), (C.c_int64 * n)()
    nb = C.c_uint64(0)
    rc = lib().orc_equity_exhaustive(h, n, board, dead, threads, w, t, l, C.byref(nb))
    return rc, list(w), list(t), list(l), nb.value


def equity_mc(holes, board, dead, seed, first, n_trials, threads=0):
    n = len(holes)
    h = (C.c_uint64 * n)(*holes)
    w, t, l = (C.c_int64 * n)(), (C.c_int64 * n)(), (C.c_int64 * n)()
    rc = lib().orc_equity_mc(h, n, board, dead, seed, first, n_trials, threads, w, t, l)
    return rc, list(w), list(t), list(l)


def _pots(pots):
    k = len(pots)
    return ((C.c_longlong * max(k, 1))(*[p[0] for p in pots]), (C.c_int * max(k, 1))(*[p[1] for p in pots]),
            (C.c_uint * max(k, 1))(*[sum(1 << s for s in p[2]) for p in pots]), k)


def equity_exhaustive_ev(holes, pots, board=0, dead=0, threads=0):
    """pots = [(bet, n, [seats])] -> (rc, win, tie, loss, chips, n_boards)"""
    n = len(holes)
    h = (C.c_uint64 * n)(*holes)
    w, t, l, ch = (C.c_int64 * n)(), (C.c_int64 * n)(), (C.c_int64 * n)(), (C.c_int64 * n)()
    nb = C.c_uint64(0)
    bet, cnt, msk, k = _pots(pots)
    L = lib()
    L.orc_equity_exhaustive_ev.argtypes = [C.POINTER(C.c_uint64), C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(C.c_longlong),
                                           C.POINTER(C.c_int), C.POINTER(C.c_uint), C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                           C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]
    rc = L.orc_equity_exhaustive_ev(h, n, board, dead, threads, bet, cnt, msk, k, w, t, l, ch, C.byref(nb))
    return rc, list(w), list(t), list(l), list(ch), nb.value


def equity_mc_ev(holes, pots, board, dead, seed, first, n_trials, threads=0):
    n = len(holes)
    h = (C.c_uint64 * n)(*holes)
    w, t, l, ch = (C.c_int64 * n)(), (C.c_int64 * n)(), (C.c_int64 * n)(), (C.c_int64 * n)()
    bet, cnt, msk, k = _pots(pots)
    L = lib()
    L.orc_equity_mc_ev.argtypes = [C.POINTER(C.c_uint64), C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int,
                                   C.POINTER(C.c_longlong), C.POINTER(C.c_int), C.POINTER(C.c_uint), C.c_int, C.POINTER(C.c_int64),
                                   C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    rc = L.orc_equity_mc_ev(h, n, board, dead, seed, first, n_trials, threads, bet, cnt, msk, k, w, t, l, ch)
    return rc, list(w), list(t), list(l), list(ch)


def equity_mc_ranges(weights, holes, board, dead, seed, first, n_trials, threads=0):
    """weights: one list of 1326 ints per seat (None for a fixed seat) -> (rc, win, tie, loss, void)"""
    import numpy as np
    n = len(holes)
    w = np.zeros((n, 1326), dtype=np.int32)
    for s in range(n):
        if weights[s] is not None:
            w[s] = np.asarray(weights[s], dtype=np.int32)
    h = (C.c_uint64 * n)(*holes)
    wi, ti, lo = (C.c_int64 * n)(), (C.c_int64 * n)(), (C.c_int64 * n)()
    void = C.c_uint64(0)
    L = lib()
    L.orc_equity_mc_ranges.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64,
                                       C.c_uint64, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                       C.POINTER(C.c_uint64)]
    rc = L.orc_equity_mc_ranges(w.ctypes.data, h, n, board, dead, seed, first, n_trials, threads, wi, ti, lo, C.byref(void))
    return rc, list(wi), list(ti), list(lo), void.value


def philox(ctr, key):
    c, k, o = (C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), (C.c_uint32 * 4)()
    lib().orc_philox4x32_10(c, k, o)
    return list(o)
