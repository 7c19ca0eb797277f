"""Where K3's fixed cost goes: the same mc_equity_exhaustive_dev launch (config 3, and over ONE board) issued (a) call by call from
Python and (b) replayed from a CUDA graph (no host work between launches), beside two reference points replayed the same way:
a kernel with no table (mc_random_hands_dev, 1 hand) and K1 over 2 hands (one block, 184 KB staged by TMA)."""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mcb200_loader  # noqa: E402

pkg = mcb200_loader.load()
E, api = pkg.engine, pkg.api
torch.cuda.set_device(0)
E.init(devices=[0])
L = E.lib()
holes = (C.c_uint64 * 2)(api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50]))
cnt = torch.zeros(27, dtype=torch.int64, device="cuda")
masks = torch.zeros(16, dtype=torch.int64, device="cuda")
keys = torch.zeros(16, dtype=torch.int32, device="cuda")
side = torch.cuda.Stream()
REPS = 200


def timed(fn, stream):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        for _ in range(5):
            fn(stream.cuda_stream)
        stream.synchronize()
        t0 = time.perf_counter()
        a.record(stream)
        for _ in range(REPS):
            fn(stream.cuda_stream)
        b.record(stream)
        host_us = 1e6 * (time.perf_counter() - t0) / REPS
        stream.synchronize()
    return 1e3 * a.elapsed_time(b) / REPS, host_us


def graphed(fn):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        fn(side.cuda_stream)                       # (module loading, attributes: outside the capture)
        side.synchronize()
        with torch.cuda.graph(g, stream=side):
            for _ in range(REPS):
                fn(torch.cuda.current_stream().cuda_stream)
    g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    g.replay()
    b.record()
    torch.cuda.synchronize()
    return 1e3 * a.elapsed_time(b) / REPS


cases = [("K3 config 3 (1,712,304 boards)", lambda st: E.check(L.mc_equity_exhaustive_dev(holes, 2, 0, 0, 0, 1712304, cnt.data_ptr(), st))),
         ("K3 over one board", lambda st: E.check(L.mc_equity_exhaustive_dev(holes, 2, 0, 0, 0, 1, cnt.data_ptr(), st))),
         ("random_hands, 1 hand (no table)", lambda st: E.check(L.mc_random_hands_dev(1, 0, 1, 7, masks.data_ptr(), st))),
         ("K1 over 2 hands (184 KB staged)", lambda st: E.check(L.mc_eval7_batch_dev(masks.data_ptr(), 2, keys.data_ptr(), st)))]
for name, fn in cases:
    dev_us, host_us = timed(fn, side)
    try:
        g_us = graphed(fn)
    except Exception as ex:          # noqa: BLE001
        g_us = float("nan")
        print("   (graph capture failed: %s)" % str(ex)[:100])
    print("%-34s per launch: %.2f us call by call (host issues one every %.2f us), %.2f us replayed from a CUDA graph" % (name, dev_us, host_us, g_us))
