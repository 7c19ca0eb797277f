#!/usr/bin/env python
"""bench.py -- headline measurement of the hand-ranking hot path (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W]           our arm (one process per GPU under torchrun)
  python bench.py --impl reference ...                          the reference's CPU algorithm (oracle port)

A step = one pass of K1 (eval7_stream) over one resident batch of 2^28 Philox-seeded 7-card hands per GPU
(2 GiB of masks in, 1 GiB of keys out: far larger than the 126 MB L2).  `value` = total 7-card evals/s over
all ranks with inputs resident in HBM, device-timed with CUDA events on the launching stream, max over
ranks.  `e2e` = the same metric through the host-buffer C-ABI call mc_eval7_batch (pinned host memory,
H2D + kernel + D2H inside the timed region).  `workloads` adds BASELINE configs 2, 3 and 4 (enumeration,
exhaustive AA-KK, 9-seat Monte Carlo with one NCCL all-reduce of the int64 counts at N > 1).
"""
import argparse
import ctypes as C
import json
import os
import re
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

SEED = 0x5EED5EED
BATCH_LOG2 = 28                  # hands per GPU per step (K1)
E2E_LOG2 = 26                    # hands per GPU per step through the host-buffer call
ALGO_BYTES_PER_EVAL = 12         # 8 B mask in + 4 B key out
# thread-level instructions per unit.  SURVEY 8d declared 64 / 70 / 140 / 1024 and allows lowering them "with a SASS
# count to justify it": these are the executed warp-instructions per unit that ncu measured for the current kernels
# (profiles/r01_v_k1_v5_ncu_full.txt, profiles/r01_z_k2_k5_ncu_full.txt), so frac = share of the 128-lane issue peak
OPS = {"eval": 46, "enum": 35, "board2": 69, "mc9": 612}
# K4 is bound by the shared-memory data pipe (one 128-byte wavefront per SM per clock): ncu counts 186 wavefronts per
# 32 trials (profiles/r01_k4_768_ncu_full.txt: 141 without bank conflicts -- 13 deck init, 63 deal steps, 42 card
# entries, 18 hash loads, flush -- plus 45 from the random D / V addresses of a warp sharing banks)
MC9_SMEM_WAVEFRONTS_PER_32_TRIALS = 186
# K2 (lockstep walk) is bound by the same pipe: ncu counts 12.9 wavefronts per 32 enumerated hands (profiles/r02_k2_lock_ncu.txt:
# D, V and the key at random addresses of a warp, the pair entry as one broadcast; a rank's share of a sharded run goes through the
# same kernel as long as it holds at least 2^22 hands)
ENUM_SMEM_WAVEFRONTS_PER_32_HANDS = 12.9
METRIC = "7-card evals/s"
# SHA-256 of the 27 config-4 counts ("w0,..,w8,t0,..,l8") by log2(trials), recorded from a 1-GPU run of this bench
# (seed 0x5EED5EED, seat 0 = A-A, eight random seats): any number of ranks must reproduce them
MC9_COUNTS_SHA256 = {34: "d5ad767cb51e8246fda2a9d0103ed6a21e30af1602176704feb75b1b3e469dab"}


K1_SOURCES = ("kernels.cu", "hashtab.cuh", "hashtab_build.h", "common.cuh", "eval.cuh")


def k1_source_hash():
    """SHA-256 of the sources the K1 kernel and its table images are built from (ties roofline.traffic to a capture)"""
    import hashlib
    h = hashlib.sha256()
    for f in K1_SOURCES:
        h.update(open(os.path.join(ROOT, "monte-carlo_b200", "csrc", f), "rb").read())
    return h.hexdigest()


def load_peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), float(p.get("sm_max_mhz", 1965.0)), "measured"
    except Exception:
        return 6650.0, 1965.0, "fallback"


class ClockSampler:
    """samples nvidia-smi during the timed region (B200_PROFILING.md clocks line)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "25"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = sorted(float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace(".", "").isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 9 for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def use_all_host_threads(oracle):
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU baseline is meant to use all the host cores"""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    oracle.lib().orc_set_threads(n)


def cpu_port_baseline(oracle, seconds_target=12.0):
    """the reference's algorithm (literal port, all host threads) on a bounded sample of the K1 workload"""
    use_all_host_threads(oracle)
    cores = oracle.lib().orc_max_threads()
    probe = oracle.random_hands(SEED, 0, 4000 * cores)
    t = time.perf_counter()
    oracle.eval_batch(probe, 2, 0)
    rate = len(probe) / (time.perf_counter() - t)
    n = max(len(probe), int(rate * seconds_target))
    masks = oracle.random_hands(SEED, 0, n)
    t = time.perf_counter()
    oracle.eval_batch(masks, 2, 0)
    dt = time.perf_counter() - t
    return n / dt, cores, n, dt


def run_reference(args):
    """--impl reference: the reference's own CPU algorithm for the path.  The Clojure original cannot run here
    (no JVM on the image), so this is the literal C port in oracle/ref_literal.c, all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle_lib as oracle
    use_all_host_threads(oracle)
    cores = oracle.lib().orc_max_threads()
    probe = oracle.random_hands(SEED, 0, 2000 * cores)
    t = time.perf_counter()
    oracle.eval_batch(probe, 2, 0)
    rate = len(probe) / (time.perf_counter() - t)
    budget = 120.0 / max(1, args.steps + args.warmup)            # whole run within a few minutes
    n = max(1000, int(rate * min(10.0, budget)))
    masks = oracle.random_hands(SEED, 0, n)
    for _ in range(args.warmup):
        oracle.eval_batch(masks, 2, 0)
    t = time.perf_counter()
    for _ in range(args.steps):
        oracle.eval_batch(masks, 2, 0)
    dt = time.perf_counter() - t
    value = n * args.steps / dt
    fast_t = time.perf_counter()
    oracle.eval_batch(masks, 0, 0)
    fast = n / (time.perf_counter() - fast_t)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64->u32", "data": "synthetic",
        "config": {"workload": "eval7_stream: Philox-seeded random 7-card hands (BASELINE config 1 generator), "
                               "bounded sample of %d hands per step" % n},
        "cpu_baseline": {"value": value, "unit": "evals/s", "cores": cores, "kind": "port",
                         "sample": "%d hands per step, literal port of hand_evaluator.clj (21 subsets x naive evaluator)" % n,
                         "fast_bitmask_oracle_evals_per_s": fast},
        "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference Clojure itself is unavailable on this box (no JVM); this is its algorithm restated in C",
    }
    print(json.dumps(line))


def packed48_dev(torch, masks):
    """int64 card masks on the device -> the 6-byte wire format of mc_eval_batch_packed48 (uint8, 6 per hand)"""
    m, h = masks.clone(), torch.zeros_like(masks)
    for k in range(7):
        low = m & -m
        h |= (torch.log2(low.double()) + 0.5).long() << (6 * k)      # (+0.5: log2 of a power of two, rounded)
        m &= m - 1
    return h.view(torch.uint8).reshape(-1, 8)[:, :6].reshape(-1)


def colex_ranks_dev(torch, masks):
    """int64 masks of 7-card hands on the device -> int32 colex ranks over card ids (mc_eval_batch_colex32's input)"""
    from math import comb
    binom = torch.tensor([[comb(x, k) for x in range(52)] for k in range(8)], dtype=torch.int64, device=masks.device)
    m, rank = masks.clone(), torch.zeros_like(masks)
    for k in range(1, 8):
        low = m & -m
        pos = (torch.log2(low.double()) + 0.5).long()
        rank += binom[k][13 * (pos >> 4) + (pos & 15) - 2]
        m &= m - 1
    return rank.to(torch.int32)


def pack27_dev(torch, ranks):
    """int32 colex ranks on the device (a multiple of 32 of them) -> the 27-bit stream of mc_eval_batch_colex27 (int32 words)"""
    r = ranks.to(torch.int64).reshape(-1, 32)
    out = torch.zeros((r.shape[0], 28), dtype=torch.int64, device=ranks.device)
    for j in range(32):
        w, sh = (27 * j) >> 5, (27 * j) & 31
        v = r[:, j] << sh
        out[:, w] |= v & 0xFFFFFFFF
        out[:, w + 1] |= v >> 32
    return out[:, :27].to(torch.int32).reshape(-1)      # (values >= 2^31 wrap to the same 32 bits)


def timed_region(torch, dist, world, fn, steps, warmup):
    """W untimed + K timed calls of fn(), CUDA events on the current stream, barrier + synchronize on both
    sides, max over ranks.  returns total milliseconds of the K steps."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def wall_region(torch, dist, world, fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t0)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def run_ours(args):
    import numpy as np
    import torch
    import mcb200_loader
    pkg = mcb200_loader.load()
    E, api = pkg.engine, pkg.api
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: libmcpoker has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    first_call = {}                              # one-time costs a caller pays once per process (reported, never timed as work)
    t0 = time.perf_counter()
    E.init(devices=[local])
    first_call["mc_init_s"] = time.perf_counter() - t0
    first_call["mc_init_note"] = ("evaluation of the table representatives on the GPU + hash-and-displace searches on the host + upload "
                                  "(the CUDA context already exists here: torch created it)")
    L = E.lib()
    stream = torch.cuda.current_stream().cuda_stream
    sharding = pkg.sharding
    hbm_peak, sm_max_mhz, peak_src = load_peaks()
    sms = torch.cuda.get_device_properties(local).multi_processor_count

    # ---- K1 (headline): resident batch, generated on the device with the config-1 generator ----------
    n = 1 << args.batch_log2
    masks = torch.empty(n, dtype=torch.int64, device="cuda")
    keys = torch.empty(n, dtype=torch.int32, device="cuda")
    E.check(L.mc_random_hands_dev(SEED, rank * n, n, 7, masks.data_ptr(), stream))
    launches = [0]

    def k1_step(fmt=E.KEY_NIBBLE):
        E.check(L.mc_eval7_batch_dev_fmt(masks.data_ptr(), n, keys.data_ptr(), fmt, stream))
        launches[0] += 1

    sampler = ClockSampler(local)
    sampler.start()                              # nvidia-smi needs ~0.1 s to deliver its first sample
    for _ in range(args.warmup):
        k1_step()
    launches[0] = 0
    ms = timed_region(torch, dist, world, k1_step, args.steps, 0)
    gpu_launches = launches[0]
    # the timed region lasts tens of milliseconds: keep the SAME launches going (untimed, uncounted) for 0.5 s so that
    # the clock / throttle samples are taken under the load that was just timed
    sampler.rows.clear()
    t_probe = time.perf_counter()
    while time.perf_counter() - t_probe < 0.5:
        k1_step()
        torch.cuda.synchronize()
    clocks = sampler.stop()
    clocks["window"] = "0.5 s of the same K1 launches issued right after the timed region (the region itself is %.0f ms)" % ms
    value = world * n * args.steps / (ms * 1e-3)
    per_launch_ms = ms / args.steps
    achieved_gbs = n * ALGO_BYTES_PER_EVAL / (per_launch_ms * 1e-3) / 1e9
    # correctness of what was timed: a slice of the keys against the oracle (outside the timed region)
    check_n = 1 << 20
    parity = None
    if rank == 0:
        import oracle_lib as oracle
        got = keys[:check_n].cpu().numpy().view(np.uint32)
        want = oracle.eval_batch(oracle.random_hands(SEED, 0, check_n), 0)
        parity = bool(np.array_equal(got, want))
        if not parity:
            raise SystemExit("K1 keys differ from the oracle: refusing to report a number")
    variants = {}
    for name, fmt in (("packed_key_bitplane", E.KEY_PACKED), ("nibble_flo_bitplane", E.KEY_NIBBLE_FLO),
                      ("nibble_rankset_table_v2", E.KEY_NIBBLE_TABLE), ("nibble_plane_hash_v5", E.KEY_NIBBLE_HASH)):
        v_ms = timed_region(torch, dist, world, lambda: k1_step(fmt), max(2, args.steps // 2), 1)
        variants[name] = world * n * max(2, args.steps // 2) / (v_ms * 1e-3)

    # ---- K1 on flop / turn hands (player->hand-value on 5 and 6 cards, hand_evaluator.clj:171): the same
    # kernel on its flop / turn table image, 8 B in + 4 B out per hand, against the same HBM roofline
    k1_stages = {}
    for n_cards, name in ((5, "k1_flop_5card"), (6, "k1_turn_6card")):
        E.check(L.mc_random_hands_dev(SEED + n_cards, rank * n, n, n_cards, masks.data_ptr(), stream))

        def stage_step():
            E.check(L.mc_eval_batch_dev(masks.data_ptr(), n, keys.data_ptr(), E.KEY_NIBBLE, n_cards, stream))
        st_steps = max(4, args.steps // 2)
        st_ms = timed_region(torch, dist, world, stage_step, st_steps, 3)
        if rank == 0:
            import oracle_lib as oracle
            got = keys[:check_n].cpu().numpy().view(np.uint32)
            if not np.array_equal(got, oracle.eval_batch(oracle.random_hands(SEED + n_cards, 0, check_n, n_cards), 0)):
                raise SystemExit("K1 %d-card keys differ from the oracle" % n_cards)
        gbs = n * ALGO_BYTES_PER_EVAL / (st_ms / st_steps * 1e-3) / 1e9
        k1_stages[name] = {"evals_per_s": world * n * st_steps / (st_ms * 1e-3), "ms": st_ms / st_steps, "hands_per_gpu": n,
                           "parity_checked": True if rank == 0 else None,
                           "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak,
                                        "kernel": "eval_stream_hash_kernel<6, Lay56, uint32_t>"}}
    E.check(L.mc_random_hands_dev(SEED, rank * n, n, 7, masks.data_ptr(), stream))       # back to the river hands (and their keys) for e2e
    k1_step()
    torch.cuda.synchronize()

    # ---- e2e: the host-buffer C-ABI calls, pinned memory, H2D + kernels + D2H inside the timed region ----
    # Three wire formats of the same batch call, all bound by the host link:
    #   colex27   3.375 B in (the same ranks as a 27-bit stream) + 2 B out     mc_eval_batch_colex27   <- the headline e2e
    #   colex32   4 B in (colex rank of the 7 cards) + 2 B out (class id)      mc_eval_batch_colex32
    #   packed48  6 B in (seven 6-bit card positions) + 4 B out (nibble key)   mc_eval_batch_packed48
    #   mask      8 B in (card mask) + 4 B out                                 mc_eval7_batch
    # The compact inputs are produced on the device here (bit shuffling only) and land in pinned host memory before the
    # timing; the run fails unless all three calls return the same values.
    ne = 1 << args.e2e_log2
    h_masks = torch.empty(ne, dtype=torch.int64).pin_memory()
    h_keys = torch.empty(ne, dtype=torch.int32).pin_memory()
    h_masks.copy_(masks[:ne])
    h_pk = torch.empty(6 * ne, dtype=torch.uint8).pin_memory()
    h_pk.copy_(packed48_dev(torch, masks[:ne]))
    h_keys48 = torch.empty(ne, dtype=torch.int32).pin_memory()
    h_ranks = torch.empty(ne, dtype=torch.int32).pin_memory()
    h_ranks.copy_(colex_ranks_dev(torch, masks[:ne]))
    h_cls = torch.empty(ne, dtype=torch.int16).pin_memory()
    h_s27 = torch.empty(27 * ne // 32, dtype=torch.int32).pin_memory()
    h_s27.copy_(pack27_dev(torch, colex_ranks_dev(torch, masks[:ne])))
    h_cls27 = torch.empty(ne, dtype=torch.int16).pin_memory()
    torch.cuda.synchronize()
    e2e_steps = max(2, args.steps // 2)

    def e2e_step():
        E.check(L.mc_eval7_batch(h_masks.data_ptr(), ne, h_keys.data_ptr()))

    def e2e48_step():
        E.check(L.mc_eval_batch_packed48(h_pk.data_ptr(), ne, h_keys48.data_ptr()))

    def e2e32_step():
        E.check(L.mc_eval_batch_colex32(h_ranks.data_ptr(), ne, h_cls.data_ptr()))

    def e2e27_step():
        E.check(L.mc_eval_batch_colex27(h_s27.data_ptr(), ne, h_cls27.data_ptr()))

    e2e_ms = wall_region(torch, dist, world, e2e_step, e2e_steps, 1)
    e2e_value = world * ne * e2e_steps / (e2e_ms * 1e-3)
    if rank == 0 and not torch.equal(h_keys[:check_n], keys[:check_n].cpu()):
        raise SystemExit("e2e keys differ from the device-resident run")
    e2e48_ms = wall_region(torch, dist, world, e2e48_step, e2e_steps, 1)
    e2e48_value = world * ne * e2e_steps / (e2e48_ms * 1e-3)
    if not torch.equal(h_keys48, h_keys):
        raise SystemExit("packed48 keys differ from the mask call")
    e2e32_ms = wall_region(torch, dist, world, e2e32_step, e2e_steps, 2)
    e2e32_value = world * ne * e2e_steps / (e2e32_ms * 1e-3)
    class_keys = torch.from_numpy(api.class_keys().astype(np.int64)).to(torch.int32)
    if not torch.equal(class_keys[h_cls.long() & 0xFFFF], h_keys):
        raise SystemExit("colex32 class ids differ from the mask call's keys")
    e2e27_ms = wall_region(torch, dist, world, e2e27_step, e2e_steps, 2)
    e2e27_value = world * ne * e2e_steps / (e2e27_ms * 1e-3)
    if not torch.equal(h_cls27, h_cls):
        raise SystemExit("colex27 class ids differ from the colex32 call's")
    # the same call on pageable buffers (what a caller gets without mc_alloc_host / pinning)
    pg_masks, pg_keys = h_masks.numpy().copy(), np.empty(ne, dtype=np.uint32)
    E.check(L.mc_eval7_batch(pg_masks.ctypes.data, ne, pg_keys.ctypes.data))
    t0 = time.perf_counter()
    E.check(L.mc_eval7_batch(pg_masks.ctypes.data, ne, pg_keys.ctypes.data))
    pageable_value = ne / (time.perf_counter() - t0)
    del pg_masks, pg_keys
    # The host link alone, measured as a ceiling for ALL ranks at once: every rank copies the headline call's pinned buffers
    # (27-bit stream in, class ids out) at the same time, in the pipeline's own 4 Mi-hand pieces on two streams, nothing
    # else running (barrier + synchronize on both sides, max over ranks): H2D alone, D2H alone, both directions together.
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    d_s27, d_cls = torch.empty(27 * ne // 32, dtype=torch.int32, device="cuda"), torch.empty(ne, dtype=torch.int16, device="cuda")
    piece = min(ne, 1 << 22)

    def copy_h2d():
        with torch.cuda.stream(s_in):
            for lo in range(0, ne, piece):
                a, b = 27 * lo // 32, 27 * (lo + piece) // 32
                d_s27[a:b].copy_(h_s27[a:b], non_blocking=True)

    def copy_d2h():
        with torch.cuda.stream(s_out):
            for lo in range(0, ne, piece):
                h_cls27[lo:lo + piece].copy_(d_cls[lo:lo + piece], non_blocking=True)

    def copy_both():
        copy_h2d()
        copy_d2h()

    reps = 4
    in_b, out_b = 27 / 8, 2
    h2d_gbs = world * in_b * ne * reps / (wall_region(torch, dist, world, copy_h2d, reps, 1) * 1e-3) / 1e9
    d2h_gbs = world * out_b * ne * reps / (wall_region(torch, dist, world, copy_d2h, reps, 1) * 1e-3) / 1e9
    both_ms = wall_region(torch, dist, world, copy_both, reps, 1)
    both_evals = world * ne * reps / (both_ms * 1e-3)               # hands/s if the copies were all there is to do
    link = {"all_ranks_at_once": True, "h2d_alone_gbs_total": h2d_gbs, "d2h_alone_gbs_total": d2h_gbs,
            "h2d_while_d2h_gbs_total": in_b * both_evals / 1e9, "d2h_while_h2d_gbs_total": out_b * both_evals / 1e9,
            "bare_copies_evals_per_s": both_evals,
            "bare_copies_note": "bare cudaMemcpyAsync of the colex27 call's buffers (3.375 B in + 2 B out per hand) in 4 Mi-hand pieces, both "
                                "directions at once, every rank at the same time, nothing else running: what the host link of this box "
                                "delivers to plain copies (the pipelined call can come out slightly above it: three streams keep more "
                                "copies queued than the two of this leg)",
            "colex27_frac_of_bare_copies": e2e27_value / both_evals,
            "colex27_h2d_gbs_total": in_b * e2e27_value / 1e9, "colex27_d2h_gbs_total": out_b * e2e27_value / 1e9,
            "colex32_h2d_gbs_total": 4 * e2e32_value / 1e9,
            "packed48_h2d_gbs_total": 6 * e2e48_value / 1e9, "mask_h2d_gbs_total": 8 * e2e_value / 1e9,
            "pageable_buffers_evals_per_s_per_gpu": pageable_value}
    del masks, keys, d_s27, d_cls

    # ---- integer pipes of this GPU, measured in this run: thread-ops per clock per SM ----------------
    pipes = None
    pb = os.path.join(ROOT, "tools", "pipebench")
    if rank == 0 and os.path.exists(pb):
        try:
            txt = subprocess.run([pb], capture_output=True, text=True, timeout=60).stdout
            pick = {"LOP3": "lop3", "IADD": "iadd", "IMAD": "imad", "IMAD.HI (mul.hi)": "imad_hi", "POPC": "popc", "SHF": "shf",
                    "LOP3+IMAD mix": "lop3_imad_mix", "LDS random (8192 words)": "lds_random_words"}
            pipes = {"unit": "thread-ops/clk/SM", "tool": "tools/pipebench.cu"}
            for label, num in re.findall(r"^(.*?)\s+([0-9.]+) thread-ops/clk/SM", txt, flags=re.M):
                if label in pick:
                    pipes[pick[label]] = float(num)
        except Exception as ex:          # the probe is a report, never a reason to lose the bench line
            pipes = {"error": str(ex)}

    # ---- other BASELINE configs --------------------------------------------------------------------
    work = dict(k1_stages)
    f_sm = (clocks.get("sm_mhz") or sm_max_mhz) * 1e6
    int_peak = sms * 128 * f_sm                                 # thread-level int ops/s at the measured clock

    out = torch.zeros(16, dtype=torch.int64, device="cuda")
    total7 = 133784560
    lo7, hi7 = sharding.shard_range(total7, rank, world)

    def enum_step():
        out.zero_()
        E.check(L.mc_enumerate7_dev(lo7, hi7 - lo7, out.data_ptr(), stream))
        if world > 1:
            dist.all_reduce(out, op=dist.ReduceOp.SUM)

    ms2 = timed_region(torch, dist, world, enum_step, 10, 3)
    hist = out[:9].cpu().tolist()
    golden7 = [23620800, 58877280, 31456080, 6466680, 5576460, 4051784, 3473184, 224848, 37444]
    rate2 = total7 * 10 / (ms2 * 1e-3)
    work["config2_enumerate_c52_7"] = {
        "hands_per_s": rate2, "ms": ms2 / 10, "histogram_equals_golden": hist == golden7, "sharded_over_ranks": world,
        "roofline": {"bound": "int_issue", "ops_per_unit": OPS["enum"], "achieved": rate2 * OPS["enum"] / 1e12,
                     "peak": world * int_peak / 1e12, "unit": "Tint-op/s", "frac": rate2 * OPS["enum"] / (world * int_peak)},
        "roofline_smem": {"bound": "shared-memory wavefronts", "wavefronts_per_32_hands": ENUM_SMEM_WAVEFRONTS_PER_32_HANDS,
                          "achieved": rate2 / 32 * ENUM_SMEM_WAVEFRONTS_PER_32_HANDS / 1e12, "peak": world * sms * f_sm / 1e12,
                          "unit": "T wavefronts/s", "frac": rate2 / 32 * ENUM_SMEM_WAVEFRONTS_PER_32_HANDS / (world * sms * f_sm)}}

    cnt = torch.zeros(27, dtype=torch.int64, device="cuda")
    holes2 = (C.c_uint64 * 2)(api.mask_of_ids([12, 25]), api.mask_of_ids([37, 50]))
    nb = 1712304

    def eqx_step():        # 18 us of work: NOT sharded (an all-reduce costs more than the job); every rank runs the whole job
        cnt.zero_()
        E.check(L.mc_equity_exhaustive_dev(holes2, 2, 0, 0, 0, nb, cnt.data_ptr(), stream))

    ms3 = timed_region(torch, dist, world, eqx_step, 20, 3)
    c3 = cnt.cpu().tolist()
    rate3 = nb * 20 / (ms3 * 1e-3)

    def eqx_back_to_back():
        E.check(L.mc_equity_exhaustive_dev(holes2, 2, 0, 0, 0, nb, cnt.data_ptr(), stream))
    ms3b = timed_region(torch, dist, world, eqx_back_to_back, 200, 5)
    work["config3_aa_vs_kk_exhaustive"] = {
        "boards_per_s": rate3, "ms": ms3 / 20, "us_per_launch_back_to_back": 1e3 * ms3b / 200,
        "counts_equal_golden": [c3[0], c3[1], c3[9]] == [1388072, 318176, 6056],
        "sharding": "none: every rank runs the whole 1.7 M-board job (replicas); it is too short to split",
        "roofline": {"bound": "int_issue", "ops_per_unit": OPS["board2"], "achieved": rate3 * OPS["board2"] / 1e12,
                     "peak": int_peak / 1e12, "unit": "Tint-op/s", "frac": rate3 * OPS["board2"] / int_peak,
                     "note": "microseconds of work: launch-bound"}}

    holes9 = (C.c_uint64 * 9)(api.mask_of_ids([12, 25]), 0, 0, 0, 0, 0, 0, 0, 0)

    def mc_run(total_trials):
        lo, hi = sharding.shard_range(total_trials, rank, world)
        cnt.zero_()
        E.check(L.mc_equity_mc_dev(holes9, 9, 0, 0, SEED, lo, hi - lo, cnt.data_ptr(), stream))
        if world > 1:
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)

    cal = 1 << 28
    ms_cal = timed_region(torch, dist, world, lambda: mc_run(cal), 1, 1)
    est_rate = cal / (ms_cal * 1e-3)
    mc_log2 = args.mc_log2
    while mc_log2 > 28 and (1 << mc_log2) / est_rate > args.mc_budget_s:
        mc_log2 -= 1
    trials = 1 << mc_log2
    sampler4 = ClockSampler(local)               # K4 runs for tenths of a second: its own clock sample
    sampler4.start()
    time.sleep(0.15)
    sampler4.rows.clear()
    ms4 = timed_region(torch, dist, world, lambda: mc_run(trials), 1, 0)
    clocks4 = sampler4.stop()
    f_sm4 = (clocks4.get("sm_mhz") or sm_max_mhz) * 1e6
    int_peak4 = sms * 128 * f_sm4
    c4 = cnt.cpu().tolist()
    rate4 = trials / (ms4 * 1e-3)
    # the same workload through the host-buffer call a JVM makes (mc_equity_mc: hole masks in, 27 host counts out,
    # wall clock around the call): inputs and outputs are a few hundred bytes, so end to end = the device rate
    e2e4_log2 = min(mc_log2, 32)
    lo4, hi4 = sharding.shard_range(1 << e2e4_log2, rank, world)
    got4 = [None]

    def mc_host_call():
        w4, t4, l4 = api.equity_mc(list(holes9), 0, 0, SEED, lo4, hi4 - lo4)
        got4[0] = torch.tensor(w4 + t4 + l4, dtype=torch.int64, device="cuda")
        if world > 1:
            dist.all_reduce(got4[0], op=dist.ReduceOp.SUM)
    ms4_host = wall_region(torch, dist, world, mc_host_call, 1, 1)
    g4 = got4[0].cpu().tolist()
    assert all(g4[s] + g4[9 + s] + g4[18 + s] == 1 << e2e4_log2 for s in range(9)), "mc_equity_mc counts do not add up"
    # G-invariance on the driver's run: trial t depends only on (seed, t), so the 27 counts of 2^34 trials must be the
    # same whatever the number of ranks; their SHA-256 is compared with the value recorded from a 1-GPU run
    import hashlib
    counts_sha = hashlib.sha256(",".join(str(v) for v in c4).encode()).hexdigest()
    golden_sha = MC9_COUNTS_SHA256.get(mc_log2)
    if rank == 0 and golden_sha and counts_sha != golden_sha:
        raise SystemExit("config 4: the 27 counts differ from the 1-GPU golden for 2^%d trials" % mc_log2)
    # EV variant: the same table with three side pots settled per trial (update-stacks gameplay.clj:104-116)
    cnt36 = torch.zeros(36, dtype=torch.int64, device="cuda")
    pots = [(100, 9, list(range(9))), (250, 6, [0, 2, 3, 5, 6, 8]), (400, 3, [0, 3, 8])]
    p_bet = (C.c_int64 * 3)(*[p[0] for p in pots]); p_n = (C.c_int32 * 3)(*[p[1] for p in pots])
    p_msk = (C.c_uint32 * 3)(*[sum(1 << q for q in p[2]) for p in pots])
    ev_trials = 1 << min(mc_log2, 32)

    def mc_ev_run():
        lo, hi = sharding.shard_range(ev_trials, rank, world)
        cnt36.zero_()
        E.check(L.mc_equity_mc_ev_dev(holes9, 9, 0, 0, SEED, lo, hi - lo, p_bet, p_n, p_msk, 3, cnt36.data_ptr(), stream))
        if world > 1:
            dist.all_reduce(cnt36, op=dist.ReduceOp.SUM)
    ms4ev = timed_region(torch, dist, world, mc_ev_run, 1, 1)
    c4ev = cnt36.cpu().tolist()
    work["config4_mc_9_seats_ev_3_pots"] = {
        "trials_per_s": ev_trials / (ms4ev * 1e-3), "trials": ev_trials, "ms": ms4ev, "pots": "9 x 100, 6 x 250, 3 x 400 chips (nested)",
        "seat0_chips_per_trial": c4ev[27] / ev_trials, "chips_paid_per_trial": sum(c4ev[27:36]) / ev_trials,
        "frac_of_plain_kernel": ev_trials / (ms4ev * 1e-3) / rate4,
        "note": "K4 EV variant: 640 threads x 96 registers (nine chip accumulators) against 768 x 80 of the plain kernel"}
    work["config4_mc_9_seats"] = {
        "counts_sha256": counts_sha, "counts_equal_1gpu_golden": (counts_sha == golden_sha) if golden_sha else None,
        "trials_per_s": rate4, "evals_per_s": 9 * rate4, "trials": trials, "log2_trials": mc_log2, "ms": ms4,
        "seat0_equity": (c4[0] + c4[9] / 2) / trials, "counts_sum_ok": all(c4[s] + c4[9 + s] + c4[18 + s] == trials for s in range(9)),
        "collective": "one NCCL all_reduce(int64 x 27)" if world > 1 else "none (1 rank)",
        "clocks": clocks4,
        "e2e": {"value": (1 << e2e4_log2) / (ms4_host * 1e-3), "unit": "trials/s", "log2_trials": e2e4_log2,
                "call": "mc_equity_mc (host arguments in, host counts out; wall clock)",
                "h2d_bytes_per_step": 72, "d2h_bytes_per_step": 216},
        "roofline": {"bound": "int_issue", "ops_per_unit": OPS["mc9"], "achieved": rate4 * OPS["mc9"] / 1e12,
                     "peak": world * int_peak4 / 1e12, "unit": "Tint-op/s", "frac": rate4 * OPS["mc9"] / (world * int_peak4)},
        "roofline_smem": {"bound": "shared-memory wavefronts", "wavefronts_per_32_trials": MC9_SMEM_WAVEFRONTS_PER_32_TRIALS,
                          "achieved": rate4 / 32 * MC9_SMEM_WAVEFRONTS_PER_32_TRIALS / 1e12, "peak": world * sms * f_sm4 / 1e12,
                          "unit": "T wavefronts/s", "frac": rate4 / 32 * MC9_SMEM_WAVEFRONTS_PER_32_TRIALS / (world * sms * f_sm4)}}

    # Monte Carlo with a weighted hole range per seat (deal spec R1): six seats on a flop, seat 0 holds A-A, five
    # seats each draw from a 200-hole weighted range; host call, wall clock (rank-local ranges of the trial index)
    if rank == 0 or world > 1:
        import random as _random
        rnd = _random.Random(7)
        all_holes = [(a, b) for b in range(52) for a in range(b)]
        rweights = [None]
        for _ in range(5):
            w = [0] * 1326
            for a, b in rnd.sample(all_holes, 200):
                w[api.hole_index(a, b)] = rnd.randint(1, 20)
            rweights.append(w)
        rholes = [api.mask_of_ids([12, 25])] + [0] * 5
        rboard = api.mask_of_ids([0, 14, 28])
        r_trials = 1 << 26
        rlo, rhi = sharding.shard_range(r_trials, rank, world)
        api.equity_mc_ranges(rweights, rholes, rboard, 0, SEED, 0, 1 << 16)
        got_r = [None]

        def ranges_call():
            w_, t_, l_, v_ = api.equity_mc_ranges(rweights, rholes, rboard, 0, SEED, rlo, rhi - rlo)
            got_r[0] = torch.tensor(w_ + t_ + l_ + [v_], dtype=torch.int64, device="cuda")
            if world > 1:
                dist.all_reduce(got_r[0], op=dist.ReduceOp.SUM)
        ms_r = wall_region(torch, dist, world, ranges_call, 1, 0)
        gr = got_r[0].cpu().tolist()
        work["mc_ranges_6_seats_flop"] = {
            "trials_per_s": r_trials / (ms_r * 1e-3), "trials": r_trials, "ms": ms_r, "void_trials": gr[-1],
            "seat0_equity": (gr[0] + gr[6] / 2) / max(1, r_trials - gr[-1]),
            "counts_sum_ok": all(gr[q] + gr[6 + q] + gr[12 + q] + gr[-1] == r_trials for q in range(6)),
            "call": "mc_equity_mc_ranges (alias tables built on the host per call; one thread per trial, holes that collide restart the trial)"}

    # the reference's own call site: one showdown of a full table through the host ABI (latency-bound by design)
    if rank == 0:
        board9 = api.mask_of_ids([0, 15, 28, 48, 5])
        holes_sd = (C.c_uint64 * 9)(*[api.mask_of_ids([6 + i, 29 + 2 * i]) for i in range(9)])
        keys_sd, grp_sd, ng_sd = (C.c_uint32 * 9)(), (C.c_int32 * 9)(), C.c_int32(0)
        for _ in range(20):
            E.check(L.mc_showdown_groups(board9, holes_sd, 9, keys_sd, grp_sd, C.byref(ng_sd), None))
        t0 = time.perf_counter()
        for _ in range(500):
            E.check(L.mc_showdown_groups(board9, holes_sd, 9, keys_sd, grp_sd, C.byref(ng_sd), None))
        work["showdown_9_seats_host_call"] = {"latency_us": 1e6 * (time.perf_counter() - t0) / 500,
                                              "note": "end-game ranking + group-winners of one table (gameplay.clj:127-132) per call; "
                                                      "seats by kernel parameter, result written to mapped pinned host memory: one launch + one sync"}
        suits1, ranks1, key1 = (C.c_int32 * 7)(0, 1, 2, 3, 0, 1, 2), (C.c_int32 * 7)(14, 14, 9, 8, 5, 3, 2), C.c_uint32(0)
        for _ in range(20):
            E.check(L.mc_eval_cards(suits1, ranks1, 7, C.byref(key1)))
        t0 = time.perf_counter()
        for _ in range(500):
            E.check(L.mc_eval_cards(suits1, ranks1, 7, C.byref(key1)))
        work["evaluator_one_hand_host_call"] = {"latency_us": 1e6 * (time.perf_counter() - t0) / 500, "key": "%06x" % key1.value,
                                                "note": "player->hand-value of one 7-card hand (hand_evaluator.clj:162-172) per call"}

    # config 5: the 169 x 169 preflop matrix, suit-canonical matchups dealt round-robin over the ranks, one all-reduce
    if not args.no_matrix:
        t0 = time.perf_counter()
        api.preflop_matrix_part(rank, world)                      # warm-up: builds and caches the matchup list
        torch.cuda.synchronize()
        first_call["first_mc_preflop_matrix_part_s"] = time.perf_counter() - t0
        first_call["first_mc_preflop_matrix_part_note"] = "includes the one-time suit-canonical matchup list (812,175 pairs x 24 permutations on the host)"
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        part, n_mu = api.preflop_matrix_part(rank, world)
        mt = torch.from_numpy(part).cuda()
        if world > 1:
            dist.all_reduce(mt, op=dist.ReduceOp.SUM)
        torch.cuda.synchronize()
        s5 = time.perf_counter() - t0
        if world > 1:
            tt = torch.tensor([s5, float(n_mu)], dtype=torch.float64, device="cuda")
            tmax = tt.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(tt, op=dist.ReduceOp.SUM)
            s5, n_mu = float(tmax[0].item()), int(tt[1].item())
        mh = mt.cpu().numpy()
        aa, kk = api.hole_class(14, 14, False), api.hole_class(13, 13, False)
        boards5 = n_mu * 1712304
        work["config5_preflop_matrix"] = {
            "seconds": s5, "matchups": n_mu, "boards_per_s": boards5 / s5, "evals_per_s": 2 * boards5 / s5,
            "cell_aa_kk_equals_golden": mh[aa, kk].tolist() == [50371344, 267444, 11004156],
            "antisymmetric": bool((mh[:, :, 0] == mh[:, :, 2].T).all()),
            "timing": "wall clock around mc_preflop_matrix_part (host call: upload of the matchup list, K5, download) + all_reduce",
            "roofline": {"bound": "int_issue", "ops_per_unit": OPS["board2"], "achieved": boards5 / s5 * OPS["board2"] / 1e12,
                         "peak": world * int_peak / 1e12, "unit": "Tint-op/s", "frac": boards5 / s5 * OPS["board2"] / (world * int_peak)}}

        # the same matrix by boards: canonical boards dealt round-robin over the ranks, all-reduce of the wins
        t0 = time.perf_counter()
        api.preflop_wins_by_boards(rank, world)                   # warm-up: builds and caches the canonical board list
        torch.cuda.synchronize()
        first_call["first_mc_preflop_wins_by_boards_s"] = time.perf_counter() - t0
        first_call["first_mc_preflop_wins_by_boards_note"] = "includes the one-time enumeration of the 134,459 suit-canonical boards on the host"
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        wins, n_b = api.preflop_wins_by_boards(rank, world)
        wt = torch.from_numpy(wins).cuda()
        if world > 1:
            dist.all_reduce(wt, op=dist.ReduceOp.SUM)
        torch.cuda.synchronize()
        mb = api.preflop_matrix_from_wins(wt.cpu().numpy())
        s5b = time.perf_counter() - t0
        if world > 1:
            tt = torch.tensor([s5b, float(n_b)], dtype=torch.float64, device="cuda")
            tmax = tt.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(tt, op=dist.ReduceOp.SUM)
            s5b, n_b = float(tmax[0].item()), int(tt[1].item())
        work["config5_preflop_matrix_by_boards"] = {
            "seconds": s5b, "canonical_boards": n_b, "hole_evals": n_b * 1081,
            "equals_brute_force_matrix": bool((mb == mh).all()),
            "timing": "wall clock around mc_preflop_wins_by_boards (K7a + K7b, download) + all_reduce + mc_preflop_matrix_from_wins",
            "note": "sort-and-count per suit-canonical board: 1,081 holes evaluated once per board instead of "
                    "2 x 1,712,304 evaluations per matchup"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle_lib as oracle
        v, cores, sample_n, dt = cpu_port_baseline(oracle)
        t0 = time.perf_counter()
        fast_n = 1 << 22
        oracle.eval_batch(oracle.random_hands(SEED, 0, fast_n), 0, 0)
        cpu = {"value": v, "unit": "evals/s", "cores": cores, "kind": "port",
               "sample": "%d of the Philox-seeded hands, literal port of hand_evaluator.clj (21 subsets x naive evaluator), %.1f s" % (sample_n, dt),
               "fast_bitmask_oracle_evals_per_s": fast_n / (time.perf_counter() - t0)}

    if rank == 0:
        # dram bytes per launch from the ncu --set full capture of THIS kernel source (tools/k1_traffic.py records the hash
        # of the K1 sources beside the counters); a capture of other sources is not reported
        traffic, traffic_note = None, "no capture (profiles/k1_traffic.json)"
        tp = os.path.join(ROOT, "profiles", "k1_traffic.json")
        if os.path.exists(tp):
            try:
                tj = json.load(open(tp))
                if tj.get("k1_source_sha256") == k1_source_hash() and args.batch_log2 == tj.get("batch_log2", 28):
                    traffic, traffic_note = tj.get("dram_bytes_per_launch_at_bench_size"), "ncu capture of these K1 sources: " + tj.get("source", "")
                else:
                    traffic_note = "the K1 sources (or the batch size) changed since the ncu capture in profiles/k1_traffic.json: not reported"
            except Exception:
                traffic = None
        line = {
            "metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": per_launch_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64->u32", "data": "synthetic",
            "config": {"workload": "eval7_stream (K1): 2^%d Philox-seeded 7-card hands per GPU resident in HBM "
                                   "(BASELINE config 1 generator at throughput size), nibble keys out" % args.batch_log2,
                       "hands_per_gpu_per_step": n, "l2": "inputs (2 GiB) + outputs (1 GiB) per step exceed the 126 MB L2",
                       "parallelism": "independent slices per rank, no data-path collective"},
            "roofline": {"bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                         "frac": achieved_gbs / hbm_peak, "traffic": traffic, "traffic_source": traffic_note, "peak_source": peak_src,
                         "kernel": "eval_stream_hash_kernel<6, Lay7, uint32_t>", "algorithmic_bytes_per_eval": ALGO_BYTES_PER_EVAL,
                         "int_issue_frac": (value / world) * OPS["eval"] / int_peak, "ops_per_eval": OPS["eval"]},
            "e2e": {"value": e2e27_value, "unit": "evals/s", "h2d_bytes_per_step": 27 * ne // 8, "d2h_bytes_per_step": 2 * ne,
                    "hands_per_step": ne,
                    "call": "mc_eval_batch_colex27 (pinned host buffers; colex ranks as a 27-bit stream in, 2-byte class ids out; 3-stage "
                            "copy / unrank / K1 / copy pipeline); values checked equal to the colex32 and mask calls' in this run",
                    "colex32_call": {"value": e2e32_value, "h2d_bytes_per_step": 4 * ne, "d2h_bytes_per_step": 2 * ne,
                                     "call": "mc_eval_batch_colex32 (4-byte colex ranks in, 2-byte class ids out)"},
                    "packed48_call": {"value": e2e48_value, "h2d_bytes_per_step": 6 * ne, "d2h_bytes_per_step": 4 * ne,
                                      "call": "mc_eval_batch_packed48 (6-byte hands in, 4-byte nibble keys out)"},
                    "mask_call": {"value": e2e_value, "h2d_bytes_per_step": 8 * ne, "d2h_bytes_per_step": 4 * ne,
                                  "call": "mc_eval7_batch (8-byte masks in, 4-byte nibble keys out)"},
                    "link": link},
            "gpu_launches": gpu_launches, "clocks": clocks, "parity_checked": parity,
            "k1_variants_evals_per_s": variants, "int_pipes_measured": pipes, "first_call_costs": first_call, "workloads": work,
        }
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch-log2", type=int, default=BATCH_LOG2)
    ap.add_argument("--e2e-log2", type=int, default=E2E_LOG2)
    ap.add_argument("--mc-log2", type=int, default=34)
    ap.add_argument("--mc-budget-s", type=float, default=20.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-matrix", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    # stdout carries exactly ONE line (the JSON).  Libraries write there too (NCCL prints its version banner to fd 1
    # on multi-rank runs), so fd 1 is pointed at stderr for the whole run and the line goes to the saved descriptor.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
