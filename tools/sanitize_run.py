"""One small call of every kernel family, for compute-sanitizer (memcheck and racecheck on the counting kernels,
which lean on shared-memory atomics, named barriers (K7b), TMA-staged tables behind an mbarrier and a mapped-host ticket
poll).  Sizes are small because the tools slow kernels down by orders of magnitude; results are still checked.
usage: compute-sanitizer --tool memcheck|racecheck python tools/sanitize_run.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import mcb200_loader  # noqa: E402
import oracle_lib as oracle  # noqa: E402

pkg = mcb200_loader.load()
E, api = pkg.engine, pkg.api
E.init(devices=[0])
masks = oracle.random_hands(1, 0, 20_000)
want = oracle.eval_batch(masks, 0)
assert np.array_equal(api.eval7_batch(masks), want)                                        # K1 river image
assert np.array_equal(api.class_keys()[api.eval_batch_colex32(api.colex_rank(masks))], want)   # unrank + K1 class flavour
assert np.array_equal(api.eval_batch_packed48(api.pack48(masks)), want)                    # unpack48
ranks = api.colex_rank(masks)
assert np.array_equal(api.class_keys()[api.eval_batch_colex27(api.pack27(ranks), len(ranks))], want)   # 27-bit stream
import torch  # noqa: E402  (device buffers of the caller for the allocation-free call)
d_r = torch.from_numpy(ranks.view(np.int32).copy()).cuda()
d_c = torch.empty(len(ranks), dtype=torch.int16, device="cuda")
E.check(E.lib().mc_eval_batch_colex32_dev_noalloc(d_r.data_ptr(), len(ranks), d_c.data_ptr(), torch.cuda.current_stream().cuda_stream))
torch.cuda.synchronize()
assert np.array_equal(api.class_keys()[d_c.cpu().numpy().view(np.uint16)], want)           # static-scratch variant
flop = oracle.random_hands(1, 0, 20_000, 5)
assert np.array_equal(api.eval7_batch(flop), oracle.eval_batch(flop, 0))                   # K1 flop / turn image
mixed = np.concatenate([masks[:500], flop[:500]])
assert np.array_equal(api.eval7_batch(mixed), oracle.eval_batch(mixed, 0))                 # slow path
hist, ks = api.enumerate7(1000, 300_000)                                                   # K2
ho, so, _ = oracle.enumerate_hands(7, 0, first=1000, count=300_000)
assert hist == ho and ks == so
aa, kk = api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50])
flop3 = api.mask_of_ids([0, 14, 28])
assert api.equity_exhaustive([aa, kk], flop3)[:3] == tuple(oracle.equity_exhaustive([aa, kk], flop3)[1:4])   # K3
pots = [(10, 2, [0, 1]), (5, 1, [1])]
assert api.equity_exhaustive_ev([aa, kk], pots, flop3)[:4] == tuple(oracle.equity_exhaustive_ev([aa, kk], pots, flop3)[1:5])
h9 = [aa] + [0] * 8
assert api.equity_mc(h9, 0, 0, 7, 0, 20_000) == tuple(oracle.equity_mc(h9, 0, 0, 7, 0, 20_000)[1:])          # K4
p9 = [(100, 9, list(range(9))), (250, 6, [0, 2, 3, 5, 6, 8]), (400, 3, [0, 3, 8])]
assert api.equity_mc_ev(h9, p9, 0, 0, 7, 0, 20_000) == tuple(oracle.equity_mc_ev(h9, p9, 0, 0, 7, 0, 20_000)[1:])   # K4 EV
w = [0] * 1326
for i in range(0, 1326, 7):
    w[i] = 1 + i % 5
assert api.equity_mc_ranges([w, w, None], [0, 0, kk], flop3, 0, 3, 0, 20_000) == tuple(
    oracle.equity_mc_ranges([w, w, None], [0, 0, kk], flop3, 0, 3, 0, 20_000)[1:])                           # range Monte Carlo
wa, wb = [0] * 1326, [0] * 1326
wa[api.hole_index(12, 25)] = 2; wa[api.hole_index(11, 24)] = 1; wb[api.hole_index(37, 50)] = 3
got, total = api.equity_ranges(wa, wb, flop3, 0)                                            # K6
assert sum(got) == total and total > 0
keys, winners, groups, ng = api.showdown(api.mask_of_ids([8, 9, 10, 11, 12]), [api.mask_of_ids([13, 14]), api.mask_of_ids([26, 27])])
assert winners == 3 and groups == [0, 0]                                                   # showdown (mapped-host ticket)
assert api.eval_cards([0, 0, 0, 0, 0], [9, 8, 7, 6, 5]) == 0x898765
cells = api.preflop_cells([(168, 154)])                                                    # K5: 3 canonical matchups
assert cells[0].tolist() == [50371344, 267444, 11004156]
wins, nb = api.preflop_wins_by_boards(0, 512)                                              # K7a + K7b on 1/512 of the boards
assert nb > 0 and int(wins.sum()) > 0
E.shutdown()
print("sanitize_run ok")
