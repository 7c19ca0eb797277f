"""ctypes binding of libmcpoker.so (include/mcpoker.h) -- the same C ABI a JVM would bind through JNA /
Panama.  No evaluation happens in Python; every compute call goes to the CUDA library and
raises if the library or a B200 is missing (there is no CPU fallback)."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MCP_LIB_PATH") or os.path.join(_HERE, "libmcpoker.so")       # (MCP_LIB_PATH: the checked build)
CSRC = os.path.join(_HERE, "csrc")

MC_OK, MC_ERR_INVALID, MC_ERR_NOT_INIT, MC_ERR_CUDA, MC_ERR_NO_DEVICE, MC_ERR_TOO_FEW = 0, -1, -2, -3, -4, -5
KEY_NIBBLE, KEY_PACKED, KEY_NIBBLE_FLO, KEY_NIBBLE_TABLE, KEY_NIBBLE_HASH, KEY_CLASS16 = 0, 1, 2, 3, 4, 5
IN_MASK64, IN_PACKED48, IN_COLEX32, IN_COLEX27 = 0, 1, 2, 3
CLASS_INVALID = 0xFFFF
N_HAND_CLASSES = 7462
NOALLOC_MAX_HANDS = 1 << 20
KEY_INVALID = 0xFFFFFFFF
MAX_SEATS = 9
N_CLASSES = 169

EXPORTS = [
    "mc_init", "mc_init_devices", "mc_shutdown", "mc_device_count", "mc_last_error", "mc_abi_version", "mc_collective",
    "mc_card_bit", "mc_key_packed_to_nibble", "mc_key_decode",
    "mc_eval_cards", "mc_eval7_batch", "mc_eval7_batch_fmt", "mc_eval_batch_packed48", "mc_eval7_batch_dev", "mc_eval7_batch_dev_fmt",
    "mc_eval_batch", "mc_eval_batch_colex32", "mc_eval_batch_colex27", "mc_eval_batch_colex32_dev_noalloc", "mc_eval_batch_dev", "mc_unrank_colex32_dev", "mc_class_keys",
    "mc_showdown", "mc_showdown_groups", "mc_settle_pots", "mc_end_game", "mc_random_hands", "mc_random_hands_dev",
    "mc_enumerate7", "mc_enumerate7_range", "mc_enumerate7_dev",
    "mc_equity_exhaustive", "mc_equity_exhaustive_range", "mc_equity_exhaustive_dev",
    "mc_equity_mc", "mc_equity_mc_dev", "mc_equity_exhaustive_ev", "mc_equity_mc_ev", "mc_equity_mc_ev_dev", "mc_equity_ranges", "mc_equity_mc_ranges", "mc_preflop_matrix", "mc_preflop_matrix_part", "mc_preflop_cells",
    "mc_preflop_boards_count", "mc_preflop_wins_by_boards", "mc_preflop_matrix_from_wins", "mc_preflop_matrix_by_boards",
    "mc_alloc_host", "mc_free_host",
]


class McError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libmcpoker error %d: %s" % (code, msg))
        self.code = code


def build(force=False):
    """Compile libmcpoker.so for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))]
    srcs.append(os.path.join(os.path.dirname(_HERE), "include", "mcpoker.h"))
    stale = (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", CSRC, "-s"] + (["-B"] if force else []))
    return LIB_PATH


_lib = None


def lib():
    """Load the shared library (never a stand-in).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise McError(MC_ERR_NO_DEVICE, "libmcpoker.so is not built (run __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    u64, u32, i64, i32, vp, sz = C.c_uint64, C.c_uint32, C.c_int64, C.c_int32, C.c_void_p, C.c_size_t
    P = C.POINTER
    L.mc_init.argtypes = [C.c_int, C.c_uint]
    L.mc_init_devices.argtypes = [P(C.c_int), C.c_int]
    L.mc_shutdown.restype = None
    L.mc_last_error.restype = C.c_char_p
    L.mc_card_bit.argtypes = [C.c_int]
    L.mc_card_bit.restype = u64
    L.mc_key_packed_to_nibble.argtypes = [u32]
    L.mc_key_packed_to_nibble.restype = u32
    L.mc_key_decode.argtypes = [u32, P(i32), P(i32), P(i32), P(i32), P(i32)]
    L.mc_eval_cards.argtypes = [P(i32), P(i32), C.c_int, P(u32)]
    L.mc_eval7_batch.argtypes = [vp, sz, vp]
    L.mc_eval7_batch_fmt.argtypes = [vp, sz, vp, C.c_int]
    L.mc_eval_batch_packed48.argtypes = [vp, sz, vp]
    L.mc_eval7_batch_dev.argtypes = [vp, sz, vp, vp]
    L.mc_eval7_batch_dev_fmt.argtypes = [vp, sz, vp, C.c_int, vp]
    L.mc_showdown.argtypes = [u64, P(u64), C.c_int, P(u32), P(u32)]
    L.mc_showdown_groups.argtypes = [u64, P(u64), C.c_int, P(u32), P(i32), P(i32), P(u32)]
    L.mc_end_game.argtypes = [u64, P(u64), C.c_int, P(i64), P(i32), P(u32), C.c_int, P(u32), P(i32), P(i32), P(u32), P(i64)]
    L.mc_eval_batch.argtypes = [vp, C.c_int, sz, vp, C.c_int, C.c_int]
    L.mc_eval_batch_colex32.argtypes = [vp, sz, vp]
    L.mc_eval_batch_colex27.argtypes = [vp, sz, vp]
    L.mc_eval_batch_colex32_dev_noalloc.argtypes = [vp, sz, vp, vp]
    L.mc_eval_batch_dev.argtypes = [vp, sz, vp, C.c_int, C.c_int, vp]
    L.mc_unrank_colex32_dev.argtypes = [vp, sz, vp, vp]
    L.mc_class_keys.argtypes = [vp, C.c_int]
    L.mc_settle_pots.argtypes = [P(i32), C.c_int, P(i64), P(i32), P(u32), C.c_int, P(i64)]
    L.mc_random_hands.argtypes = [u64, u64, sz, C.c_int, vp]
    L.mc_random_hands_dev.argtypes = [u64, u64, sz, C.c_int, vp, vp]
    L.mc_enumerate7.argtypes = [P(u64), P(u64)]
    L.mc_enumerate7_range.argtypes = [u64, u64, P(u64), P(u64)]
    L.mc_enumerate7_dev.argtypes = [u64, u64, vp, vp]
    L.mc_equity_exhaustive.argtypes = [P(u64), C.c_int, u64, u64, P(i64), P(i64), P(i64), P(u64)]
    L.mc_equity_exhaustive_range.argtypes = [P(u64), C.c_int, u64, u64, u64, u64, P(i64), P(i64), P(i64)]
    L.mc_equity_exhaustive_dev.argtypes = [P(u64), C.c_int, u64, u64, u64, u64, vp, vp]
    L.mc_equity_mc.argtypes = [P(u64), C.c_int, u64, u64, u64, u64, u64, P(i64), P(i64), P(i64)]
    L.mc_equity_mc_dev.argtypes = [P(u64), C.c_int, u64, u64, u64, u64, u64, vp, vp]
    L.mc_equity_exhaustive_ev.argtypes = [P(u64), C.c_int, u64, u64, P(i64), P(i32), P(u32), C.c_int, P(i64), P(i64), P(i64), P(i64), P(u64)]
    L.mc_equity_mc_ev.argtypes = [P(u64), C.c_int, u64, u64, u64, u64, u64, P(i64), P(i32), P(u32), C.c_int, P(i64), P(i64), P(i64), P(i64)]
    L.mc_equity_mc_ev_dev.argtypes = [P(u64), C.c_int, u64, u64, u64, u64, u64, P(i64), P(i32), P(u32), C.c_int, vp, vp]
    L.mc_equity_mc_ranges.argtypes = [vp, P(u64), C.c_int, u64, u64, u64, u64, u64, P(i64), P(i64), P(i64), P(u64)]
    L.mc_equity_ranges.argtypes = [P(i32), P(i32), u64, u64, P(i64), P(i64)]
    L.mc_preflop_matrix.argtypes = [vp]
    L.mc_preflop_matrix_part.argtypes = [C.c_int, C.c_int, vp, P(i32)]
    L.mc_preflop_cells.argtypes = [P(i32), P(i32), C.c_int, vp]
    L.mc_preflop_boards_count.argtypes = [P(i32), P(i64)]
    L.mc_preflop_wins_by_boards.argtypes = [C.c_int, C.c_int, vp, P(i32)]
    L.mc_preflop_matrix_from_wins.argtypes = [vp, vp]
    L.mc_preflop_matrix_by_boards.argtypes = [vp]
    L.mc_alloc_host.argtypes = [C.c_size_t]
    L.mc_alloc_host.restype = C.c_void_p
    L.mc_free_host.argtypes = [C.c_void_p]
    _lib = L
    return L


def check(rc):
    if rc < 0:
        raise McError(rc, lib().mc_last_error().decode("utf-8", "replace"))
    return rc


_initialised = False


INIT_NCCL = 1


def init(n_gpus=0, devices=None, flags=0):
    global _initialised
    L = lib()
    if devices is not None:
        arr = (C.c_int * len(devices))(*devices)
        check(L.mc_init_devices(arr, len(devices)))
    else:
        check(L.mc_init(n_gpus, flags))
    _initialised = True
    return L.mc_device_count()


def ensure_init():
    if not _initialised:
        init()


def shutdown():
    global _initialised
    if _lib is not None:
        _lib.mc_shutdown()
    _initialised = False
