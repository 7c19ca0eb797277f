"""A/B of the one collective of the path inside libmcpoker.so (single process driving every device of the box, the
JVM-host mode): the int64 counts of the devices summed on the host from pinned copies (default) against one
ncclAllReduce per device in one group (mc_init(0, MC_INIT_NCCL)).  Times mc_equity_mc, 9 seats, for a tiny job (the
call is then all fixed cost: launches, copies / collective, synchronisation) and for 2^32 trials; counts must be equal.
usage: python tools/collective_ab.py        (needs >= 2 devices)"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mcb200_loader  # noqa: E402

pkg = mcb200_loader.load()
E, api = pkg.engine, pkg.api
holes = [api.mask_of_ids([12, 25])] + [0] * 8
res = {}
for name, flags in (("host_sum", 0), ("nccl_allreduce", E.INIT_NCCL)):
    t0 = time.perf_counter()
    nd = E.init(flags=flags)
    t_init = time.perf_counter() - t0
    for _ in range(20):
        api.equity_mc(holes, 0, 0, 1, 0, 1 << 12)
    t0 = time.perf_counter()
    for _ in range(200):
        small = api.equity_mc(holes, 0, 0, 1, 0, 1 << 12)
    us_small = 1e6 * (time.perf_counter() - t0) / 200
    api.equity_mc(holes, 0, 0, 1, 0, 1 << 28)
    t0 = time.perf_counter()
    big = api.equity_mc(holes, 0, 0, 1, 0, 1 << 32)
    s_big = time.perf_counter() - t0
    res[name] = (small, big)
    print("%-15s %d devices, collective=%d: init %.2f s, mc_equity_mc(2^12 trials) %.1f us per call, 2^32 trials %.1f ms = %.1f G trials/s" % (
        name, nd, E.lib().mc_collective(), t_init, us_small, 1e3 * s_big, (1 << 32) / s_big / 1e9))
print("counts equal:", res["host_sum"] == res["nccl_allreduce"])
