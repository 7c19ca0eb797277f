"""A/B of the pipeline stage size of the host-buffer calls (MCP_STAGE_LOG2=12..22): 2^26 hands, pinned buffers."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mcb200_loader  # noqa: E402

pkg = mcb200_loader.load()
E = pkg.engine
torch.cuda.set_device(0)
ndev = E.init() if os.environ.get("MCP_ALL_DEVICES") else E.init(devices=[0])      # all devices in ONE process: the JVM-host mode
L = E.lib()
n = 1 << int(os.environ.get("MCP_E2E_LOG2", "26"))
st = torch.cuda.current_stream().cuda_stream
d = torch.empty(n, dtype=torch.int64, device="cuda")
E.check(L.mc_random_hands_dev(0x5EED5EED, 0, n, 7, d.data_ptr(), st))
h_masks = torch.empty(n, dtype=torch.int64).pin_memory()
h_masks.copy_(d)
import bench  # noqa: E402
h_pk = torch.empty(6 * n, dtype=torch.uint8).pin_memory()
h_pk.copy_(bench.packed48_dev(torch, d))
h_rk = torch.empty(n, dtype=torch.int32).pin_memory()
h_rk.copy_(bench.colex_ranks_dev(torch, d))
c3 = torch.empty(n, dtype=torch.int16).pin_memory()
h_27 = torch.empty(27 * n // 32, dtype=torch.int32).pin_memory()
h_27.copy_(bench.pack27_dev(torch, bench.colex_ranks_dev(torch, d)))
c4 = torch.empty(n, dtype=torch.int16).pin_memory()
k1 = torch.empty(n, dtype=torch.int32).pin_memory()
k2 = torch.empty(n, dtype=torch.int32).pin_memory()
torch.cuda.synchronize()
res = {}
for name, fn in (("mask", lambda: E.check(L.mc_eval7_batch(h_masks.data_ptr(), n, k1.data_ptr()))),
                 ("packed48", lambda: E.check(L.mc_eval_batch_packed48(h_pk.data_ptr(), n, k2.data_ptr()))),
                 ("colex32", lambda: E.check(L.mc_eval_batch_colex32(h_rk.data_ptr(), n, c3.data_ptr()))),
                 ("colex27", lambda: E.check(L.mc_eval_batch_colex27(h_27.data_ptr(), n, c4.data_ptr())))):
    fn()
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    res[name] = n / best / 1e9
print("%d device(s) in one process, 2^%d hands, stage 2^%s: mask %.2f, packed48 %.2f, colex32 %.2f, colex27 %.2f G evals/s, equal %s" % (
    ndev, n.bit_length() - 1, os.environ.get("MCP_STAGE_LOG2", "22"), res["mask"], res["packed48"], res["colex32"], res["colex27"],
    torch.equal(k1, k2) and torch.equal(c3, c4)))
