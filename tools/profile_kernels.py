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
n = 1 << log2
masks = torch.empty(n, dtype=torch.int64, device="cuda")
keys = torch.empty(n, dtype=torch.int32, device="cuda")
E.check(L.mc_random_hands_dev(0x5EED5EED, 0, n, 7, masks.data_ptr(), st))
for rep in range(2):
    for fmt in (E.KEY_NIBBLE_HASH, E.KEY_NIBBLE_TABLE, E.KEY_PACKED):
        E.check(L.mc_eval7_batch_dev_fmt(masks.data_ptr(), n, keys.data_ptr(), fmt, st))
out = torch.zeros(16, dtype=torch.int64, device="cuda")
if which == "k1":
    torch.cuda.synchronize()
    print("ok k1")
    sys.exit(0)
for rep in range(2):
    E.check(L.mc_enumerate7_dev(0, 133784560, out.data_ptr(), st))
cnt = torch.zeros(27, dtype=torch.int64, device="cuda")
holes2 = (C.c_uint64 * 2)(api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50]))
holes9 = (C.c_uint64 * 9)(api.mask_of_ids([12, 25]), 0, 0, 0, 0, 0, 0, 0, 0)
for rep in range(2):
    E.check(L.mc_equity_exhaustive_dev(holes2, 2, 0, 0, 0, 1712304, cnt.data_ptr(), st))
    E.check(L.mc_equity_mc_dev(holes9, 9, 0, 0, 0x5EED5EED, 0, 1 << 26, cnt.data_ptr(), st))
aa, kk = api.hole_class(14, 14, False), api.hole_class(13, 13, False)
for rep in range(2):
    cell = api.preflop_cells([(aa, kk)])          # 3 suit-canonical matchups x 1,712,304 boards
torch.cuda.synchronize()
print("ok", out[:9].tolist(), cell.tolist())
