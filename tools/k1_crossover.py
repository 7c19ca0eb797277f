"""Device-resident K1 time per call by batch size: the 64 KB rank-set-table kernel (key_format 3) against the
plane-hash kernel (key_format 4), to place K1_HASH_MIN of capi.cu.  CUDA events, 200 back-to-back launches."""
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
nmax = 1 << 24
masks = torch.empty(nmax, dtype=torch.int64, device="cuda")
keys = torch.empty(nmax, dtype=torch.int32, device="cuda")
E.check(L.mc_random_hands_dev(0x5EED5EED, 0, nmax, 7, masks.data_ptr(), st))
for log2 in range(10, 25):
    n = 1 << log2
    row = []
    for fmt in (E.KEY_NIBBLE_TABLE, E.KEY_NIBBLE_HASH):
        for _ in range(20):
            E.check(L.mc_eval7_batch_dev_fmt(masks.data_ptr(), n, keys.data_ptr(), fmt, st))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(200):
            E.check(L.mc_eval7_batch_dev_fmt(masks.data_ptr(), n, keys.data_ptr(), fmt, st))
        b.record()
        torch.cuda.synchronize()
        row.append(a.elapsed_time(b) * 1e3 / 200)
    print("2^%d hands: table %.2f us  plane-hash %.2f us" % (log2, row[0], row[1]))
