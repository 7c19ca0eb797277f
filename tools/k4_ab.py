"""A/B of the K4 launch shapes (MCP_K4_THREADS=512|768|1024): 9 seats (1 fixed + 8 random), 2^log2 trials, CUDA events.
usage: MCP_K4_THREADS=768 python tools/k4_ab.py [log2_trials]   -- prints G trials/s and the 27 counts' checksum"""
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
log2 = int(sys.argv[1]) if len(sys.argv) > 1 else 32
holes9 = (C.c_uint64 * 9)(api.mask_of_ids([12, 25]), 0, 0, 0, 0, 0, 0, 0, 0)
cnt = torch.zeros(27, dtype=torch.int64, device="cuda")
E.check(L.mc_equity_mc_dev(holes9, 9, 0, 0, 0x5EED5EED, 0, 1 << 26, cnt.data_ptr(), st))
best = 1e9
for rep in range(3):
    cnt.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    E.check(L.mc_equity_mc_dev(holes9, 9, 0, 0, 0x5EED5EED, 0, 1 << log2, cnt.data_ptr(), st))
    b.record()
    torch.cuda.synchronize()
    best = min(best, a.elapsed_time(b))
c = cnt.tolist()
# the EV variant (three nested side pots), MCP_K4_EV_THREADS=640|768
cnt36 = torch.zeros(36, dtype=torch.int64, device="cuda")
p_bet, p_n, p_msk = (C.c_int64 * 3)(100, 250, 400), (C.c_int32 * 3)(9, 6, 3), (C.c_uint32 * 3)(0x1FF, 0b101101101, 0b100001001)
ev_best = 1e9
for rep in range(3):
    cnt36.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    E.check(L.mc_equity_mc_ev_dev(holes9, 9, 0, 0, 0x5EED5EED, 0, 1 << log2, p_bet, p_n, p_msk, 3, cnt36.data_ptr(), st))
    b.record()
    torch.cuda.synchronize()
    ev_best = min(ev_best, a.elapsed_time(b))
c36 = cnt36.tolist()
print("K4 EV threads=%s: %.2f G trials/s, counts equal to the plain kernel %s, chips per trial %.3f" % (
    os.environ.get("MCP_K4_EV_THREADS", "768 (default)"), (1 << log2) / ev_best / 1e6, c36[:27] == c, sum(c36[27:]) / (1 << log2)))
print("K4 threads=%s: %.2f G trials/s (%.1f ms for 2^%d), win0=%d checksum=%d" % (
    os.environ.get("MCP_K4_THREADS", "768 (default)"), (1 << log2) / best / 1e6, best, log2, c[0], sum((i + 1) * v for i, v in enumerate(c))))
