"""One launch of each hot kernel at a moderate size, for `ncu --set full` (B200_PROFILING.md).  Not a bench:
numbers printed under a profiler are never reported."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mcb200_loader  # noqa: E402

pkg = mcb200_loader.load()
E, api = pkg.engine, pkg.api
torch.cuda.set_device(0)
E.init(devices=[0])
L = E.lib()
st = torch.cuda.current_stream().cuda_stream
log2 = int(sys.argv[1]) if len(sys.argv) > 1 else 26
which = sys.argv[2] if len(sys.argv) > 2 else "all"
REPS = int(os.environ.get("MCP_PROFILE_REPS", "2"))
n = 1 << log2
masks = torch.empty(n, dtype=torch.int64, device="cuda")
keys = torch.empty(n, dtype=torch.int32, device="cuda")
E.check(L.mc_random_hands_dev(0x5EED5EED, 0, n, 7, masks.data_ptr(), st))
cls = torch.empty(n, dtype=torch.int16, device="cuda")
ranks = torch.empty(n, dtype=torch.int32, device="cuda")
for rep in range(REPS):
    E.check(L.mc_eval7_batch_dev_fmt(masks.data_ptr(), n, keys.data_ptr(), E.KEY_NIBBLE_HASH, st))         # K1, river image, keys
    E.check(L.mc_eval_batch_dev(masks.data_ptr(), n, cls.data_ptr(), E.KEY_CLASS16, 7, st))                # K1, river image, class ids
if which == "k1":
    torch.cuda.synchronize()
    print("ok k1")
    sys.exit(0)
import bench  # noqa: E402
ranks.copy_(bench.colex_ranks_dev(torch, masks))
m5 = torch.empty(n, dtype=torch.int64, device="cuda")
E.check(L.mc_random_hands_dev(0x5EED5EED + 5, 0, n, 5, m5.data_ptr(), st))
for rep in range(REPS):
    E.check(L.mc_eval_batch_dev(m5.data_ptr(), n, keys.data_ptr(), E.KEY_NIBBLE, 5, st))                   # K1, flop / turn image
    E.check(L.mc_unrank_colex32_dev(ranks.data_ptr(), n, m5.data_ptr(), st))                               # colex32 -> masks
out = torch.zeros(16, dtype=torch.int64, device="cuda")
for rep in range(REPS):
    E.check(L.mc_enumerate7_dev(0, 133784560, out.data_ptr(), st))
cnt = torch.zeros(36, dtype=torch.int64, device="cuda")
holes2 = (C.c_uint64 * 2)(api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50]))
holes9 = (C.c_uint64 * 9)(api.mask_of_ids([12, 25]), 0, 0, 0, 0, 0, 0, 0, 0)
p_bet, p_n, p_msk = (C.c_int64 * 3)(100, 250, 400), (C.c_int32 * 3)(9, 6, 3), (C.c_uint32 * 3)(0x1FF, 0b101101101, 0b100001001)
for rep in range(REPS):
    E.check(L.mc_equity_exhaustive_dev(holes2, 2, 0, 0, 0, 1712304, cnt.data_ptr(), st))
    E.check(L.mc_equity_mc_dev(holes9, 9, 0, 0, 0x5EED5EED, 0, 1 << 26, cnt.data_ptr(), st))
    E.check(L.mc_equity_mc_ev_dev(holes9, 9, 0, 0, 0x5EED5EED, 0, 1 << 26, p_bet, p_n, p_msk, 3, cnt.data_ptr(), st))
aa, kk = api.hole_class(14, 14, False), api.hole_class(13, 13, False)
for rep in range(REPS):
    cell = api.preflop_cells([(aa, kk)])          # 3 suit-canonical matchups x 1,712,304 boards
w = [1] * 1326
rng = api.equity_mc_ranges([w, w, w], [0, 0, 0], 0, 0, 5, 0, 1 << 22)     # range Monte Carlo, three uniform ranges
torch.cuda.synchronize()
print("ok", out[:9].tolist(), cell.tolist(), rng[0])
