#!/bin/bash
# One GPU-box session of evidence for profiles/: the checked build, the default bench + reference arm, the ncu launch list
# of the bench command, and ncu --set full summaries (text; the .ncu-rep files are deleted on the box: gpurun_out/ is
# capped at 64 MiB).  usage (from the repo root on the box): bash tools/gpu_profile_round.sh <tag>
T=${1:-r02}
O=gpurun_out
mkdir -p $O
MCP_LIB_PATH=$PWD/monte-carlo_b200/libmcpoker_check.so python tools/sanitize_run.py > $O/${T}_checked_build_run.txt 2>&1; tail -1 $O/${T}_checked_build_run.txt
python bench.py > $O/${T}_bench_default.json 2> $O/${T}_bench_default.err; echo "bench rc=$?"; tail -2 $O/${T}_bench_default.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_reference.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_bench_launches.csv python bench.py --steps 3 --warmup 3 --no-matrix --no-cpu > $O/ncu_launch.log 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:eval_stream_hash -c 4 -o /tmp/${T}_k1 python tools/profile_kernels.py 28 k1 > $O/ncu_k1.log 2>&1; echo "k1 rc=$?"
python tools/ncu_summary.py /tmp/${T}_k1.ncu-rep eval_stream_hash=268435456 > $O/${T}_k1_ncu_full.txt 2>&1
python tools/k1_traffic.py /tmp/${T}_k1.ncu-rep 28 > $O/${T}_k1_traffic.log 2>&1 && cp profiles/k1_traffic.json $O/k1_traffic.json
ncu -i /tmp/${T}_k1.ncu-rep --page raw --csv > $O/${T}_k1_raw.csv 2>/dev/null
MCP_PROFILE_REPS=1 ncu --set full --clock-control none -o /tmp/${T}_all python tools/profile_kernels.py 26 > $O/ncu_all.log 2>&1; echo "all rc=$?"
python tools/ncu_summary.py /tmp/${T}_all.ncu-rep eval_stream_hash=67108864 unrank=67108864 enum7=133784560 equity_exh=1712304 equity_mc_fy=67108864 preflop_hash=5136912 equity_mc_ranges=4194304 > $O/${T}_all_ncu_full.txt 2>&1
ncu -i /tmp/${T}_all.ncu-rep --page raw --csv > $O/${T}_all_raw.csv 2>/dev/null
ls -la /tmp/*.ncu-rep; du -sh $O
