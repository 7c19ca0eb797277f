"""Times K2 (all C(52,7) hands) with CUDA events and checks the histogram + key checksum against the goldens."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mcb200_loader  # noqa: E402

pkg = mcb200_loader.load()
E = pkg.engine
torch.cuda.set_device(0)
E.init(devices=[0])
L = E.lib()
st = torch.cuda.current_stream().cuda_stream
out = torch.zeros(16, dtype=torch.int64, device="cuda")
best = 1e9
for rep in range(20):
    out.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    E.check(L.mc_enumerate7_dev(0, 133784560, out.data_ptr(), st))
    b.record()
    torch.cuda.synchronize()
    best = min(best, a.elapsed_time(b))
o = out.tolist()
GOLD = [23620800, 58877280, 31456080, 6466680, 5576460, 4051784, 3473184, 224848, 37444]
print("K2 (%s walk): %.4f ms = %.1f G hands/s, golden histogram %s, key sum %s" % (
    {"w": "warp-cooperative", "t": "per-thread", "p": "pair"}.get(os.environ.get("MCP_K2_VARIANT", "")[:1], "lockstep"), best, 133784560 / best / 1e6, o[:9] == GOLD, o[9] == 307406628956200))
