"""One mc_preflop_matrix_by_boards call (K7a / K7b launches) for ncu."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcb200_loader  # noqa: E402

pkg = mcb200_loader.load()
pkg.engine.init(devices=[0])
api = pkg.api
api.preflop_boards_count()
for _ in range(2):
    t = time.perf_counter()
    m = api.preflop_matrix_by_boards()
    print("by boards: %.1f ms" % (1e3 * (time.perf_counter() - t)))
print(m[168, 154].tolist())
