"""Times K3 on BASELINE config 3 (A-A vs K-K, all C(48,5) boards) with CUDA events, golden counts checked, and the
fixed cost of a launch (one board; one board per thread)."""
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
cnt = torch.zeros(27, dtype=torch.int64, device="cuda")
holes = (C.c_uint64 * 2)(api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50]))
for _ in range(5):
    E.check(L.mc_equity_exhaustive_dev(holes, 2, 0, 0, 0, 1712304, cnt.data_ptr(), st))
cnt.zero_()
reps = 200
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(reps):
    E.check(L.mc_equity_exhaustive_dev(holes, 2, 0, 0, 0, 1712304, cnt.data_ptr(), st))
b.record()
torch.cuda.synchronize()
c = cnt.tolist()
ok = (c[0], c[1], c[9]) == (1388072 * reps, 318176 * reps, 6056 * reps)
print("K3: %.2f us per launch (back to back), golden counts %s" % (1e3 * a.elapsed_time(b) / reps, ok))
# the fixed cost of a launch: the same kernel over ONE board (launch + table staging + counter flush)
a.record()
for _ in range(reps):
    E.check(L.mc_equity_exhaustive_dev(holes, 2, 0, 0, 0, 1, cnt.data_ptr(), st))
b.record()
torch.cuda.synchronize()
print("K3 over one board (launch + staging floor, 1 block): %.2f us" % (1e3 * a.elapsed_time(b) / reps))
for nb in (151552, 151552 * 4):
    a.record()
    for _ in range(reps):
        E.check(L.mc_equity_exhaustive_dev(holes, 2, 0, 0, 0, nb, cnt.data_ptr(), st))
    b.record()
    torch.cuda.synchronize()
    print("K3 over %d boards (148 blocks): %.2f us" % (nb, 1e3 * a.elapsed_time(b) / reps))
