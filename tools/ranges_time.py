"""Times mc_equity_mc_ranges (deal spec R1) on the bench's range workload (six seats on a flop: A-A fixed, five 200-hole weighted ranges)
and on three uniform 1326-hole ranges preflop; wall clock around the host call."""
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcb200_loader  # noqa: E402

pkg = mcb200_loader.load()
E, api = pkg.engine, pkg.api
E.init(devices=[0])
rnd = random.Random(7)
holes = [(a, b) for b in range(52) for a in range(b)]
weights = [None]
for _ in range(5):
    w = [0] * 1326
    for a, b in rnd.sample(holes, 200):
        w[api.hole_index(a, b)] = rnd.randint(1, 20)
    weights.append(w)
cases = [("6 seats, flop, five 200-hole ranges", weights, [api.mask_of_ids([12, 25])] + [0] * 5, api.mask_of_ids([0, 14, 28])),
         ("3 seats, preflop, uniform ranges", [[1] * 1326] * 3, [0, 0, 0], 0)]
for name, wts, hm, board in cases:
    api.equity_mc_ranges(wts, hm, board, 0, 5, 0, 1 << 16)
    n = 1 << 26
    t0 = time.perf_counter()
    w, t, l, void = api.equity_mc_ranges(wts, hm, board, 0, 5, 0, n)
    dt = time.perf_counter() - t0
    print("%s: %.2f G trials/s (%.1f ms for 2^26), seat-0 equity %.4f, void %d, checksum %d" % (
        name, n / dt / 1e9, 1e3 * dt, (w[0] + t[0] / 2) / n, void, sum((i + 1) * v for i, v in enumerate(w + t + l))))
