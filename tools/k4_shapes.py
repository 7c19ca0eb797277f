"""K4 rate by table shape: seat 0 holds A-A, the other seats and the board are random; 2^31 trials each, CUDA events."""
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
n = 1 << 31
for seats, board in ((2, []), (3, []), (6, []), (9, []), (2, [5, 6, 7]), (6, [5, 6, 7]), (9, [5, 6, 7])):
    holes = (C.c_uint64 * 9)(*([api.mask_of_ids([12, 25])] + [0] * 8))
    bm = api.mask_of_ids(board) if board else 0
    E.check(L.mc_equity_mc_dev(holes, seats, bm, 0, 1, 0, 1 << 24, cnt.data_ptr(), st))
    best = 1e9
    for _ in range(3):
        cnt.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        E.check(L.mc_equity_mc_dev(holes, seats, bm, 0, 1, 0, n, cnt.data_ptr(), st))
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    c = cnt.tolist()
    print("%d seats, %s: %6.1f G trials/s = %6.1f G evals/s, seat 0 equity %.4f" % (
        seats, "flop known" if board else "preflop   ", n / best / 1e6, seats * n / best / 1e6, (c[0] + c[9] / 2) / n))
