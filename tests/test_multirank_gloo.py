"""CPU-only, world_size 2 over gloo: the N > 1 host logic (range sharding + one all-reduce of int64 counts).
The per-rank compute is stood in for by the CPU oracle -- this test checks the plumbing, not the kernels."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, results):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import mcb200_loader
    import oracle_lib as oracle
    sh = mcb200_loader.load().sharding
    holes = [oracle.card_bit(12) | oracle.card_bit(25)] + [0] * 8
    # Monte Carlo: contiguous Philox counter ranges
    n_trials = 40_001
    lo, hi = sh.shard_range(n_trials, rank, world)
    rc, w, t, l = oracle.equity_mc(holes, 0, 0, 0x5EED5EED, lo, hi - lo, threads=2)
    counts = torch.tensor(w + t + l, dtype=torch.int64)
    sh.allreduce_counts(counts, dist)
    # enumeration: contiguous combination-index ranges (6-card hands for speed)
    total6 = 20358520
    lo, hi = sh.shard_range(total6, rank, world)
    h, ks, _ = oracle.enumerate_hands(6, 0, threads=2, first=lo, count=hi - lo)
    hist = torch.tensor(h + [ks], dtype=torch.int64)
    sh.allreduce_counts(hist, dist)
    if rank == 0:
        results["mc"] = counts.tolist()
        results["enum"] = hist.tolist()
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_equals_single_rank(oracle):
    world, port = 2, 29500 + os.getpid() % 2000
    with mp.Manager() as mgr:
        results = mgr.dict()
        mp.spawn(_worker, args=(world, port, results), nprocs=world, join=True)
        holes = [oracle.card_bit(12) | oracle.card_bit(25)] + [0] * 8
        rc, w, t, l = oracle.equity_mc(holes, 0, 0, 0x5EED5EED, 0, 40_001)
        assert results["mc"] == w + t + l
        h, ks, _ = oracle.enumerate_hands(6, 0)
        assert results["enum"] == h + [ks]


def test_shard_range_tiles_exactly(pkg):
    sh = pkg.sharding
    for total in (0, 1, 7, 133784560, 1 << 34):
        for world in (1, 2, 3, 4, 8):
            edges = [sh.shard_range(total, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sh.shard_range(10, 2, 2)


# ---- the same plumbing with the REAL per-rank compute: one process per GPU, libmcpoker per rank ------------------
def _gpu_worker(rank, world, port, results):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    ndev = torch.cuda.device_count()
    dev = rank % ndev                                   # (on a 1-GPU box both ranks share device 0: the plumbing is the same)
    torch.cuda.set_device(dev)
    use_nccl = ndev >= world
    dist.init_process_group("nccl" if use_nccl else "gloo", rank=rank, world_size=world,
                            **({"device_id": torch.device("cuda", dev)} if use_nccl else {}))
    import mcb200_loader
    pkg = mcb200_loader.load()
    pkg.engine.init(devices=[dev])                      # mc_init_devices({LOCAL_RANK}), as bench.py does under torchrun
    api, sh = pkg.api, pkg.sharding
    where = "cuda" if use_nccl else "cpu"
    holes = [api.mask_of_ids([12, 25])] + [0] * 8
    n_trials = 3_000_001
    lo, hi = sh.shard_range(n_trials, rank, world)
    w, t, l = api.equity_mc(holes, 0, 0, 0x5EED5EED, lo, hi - lo)
    counts = torch.tensor(w + t + l, dtype=torch.int64, device=where)
    sh.allreduce_counts(counts, dist)
    total7 = 133784560
    lo, hi = sh.shard_range(total7, rank, world)
    h, ks = api.enumerate7(lo, hi - lo)
    hist = torch.tensor(h + [ks], dtype=torch.int64, device=where)      # (the key checksum is 3.07e14: far inside int64)
    sh.allreduce_counts(hist, dist)
    if rank == 0:
        results["mc"] = counts.cpu().tolist()
        results["enum"] = hist.cpu().tolist()
        results["backend"] = "nccl" if use_nccl else "gloo"
    dist.barrier()
    pkg.engine.shutdown()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_two_ranks_on_gpus_equal_one_rank(gpu, oracle):
    """two processes, each driving its own device through mc_init_devices (NCCL all-reduce of the counts when the box has
    two GPUs, gloo otherwise): Monte Carlo counts == the single-process call == the oracle; enumeration == the goldens"""
    world, port = 2, 31500 + os.getpid() % 2000
    holes = [gpu.api.mask_of_ids([12, 25])] + [0] * 8
    single = gpu.api.equity_mc(holes, 0, 0, 0x5EED5EED, 0, 3_000_001)
    with mp.Manager() as mgr:
        results = mgr.dict()
        mp.spawn(_gpu_worker, args=(world, port, results), nprocs=world, join=True)
        assert results["mc"] == single[0] + single[1] + single[2]
        assert results["enum"][:9] == [23620800, 58877280, 31456080, 6466680, 5576460, 4051784, 3473184, 224848, 37444]
        assert results["enum"][9] == 307406628956200
    rc, w, t, l = oracle.equity_mc(holes, 0, 0, 0x5EED5EED, 0, 200_000)
    assert gpu.api.equity_mc(holes, 0, 0, 0x5EED5EED, 0, 200_000) == (w, t, l)
