"""SASS evidence for profiles/: per hot kernel of libmcpoker.so the opcode histogram and every memory / TMA / barrier /
wide-multiply instruction (the lines DESIGN.md argues from: UBLKCP = TMA bulk copy, SYNCS = mbarrier, LDG.E.*.128 streaming
loads, LDS.U16 table loads by 32-bit address, IMAD.WIDE / IMAD.HI, POPC, VIMNMX3, REDUX, ATOMS).  Runs without a GPU.
usage: python tools/sass_excerpts.py [out_dir=profiles] [tag=r02]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "monte-carlo_b200", "libmcpoker.so")
out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles")
tag = sys.argv[2] if len(sys.argv) > 2 else "r02"
KERNELS = [("k1_river_keys", r"eval_stream_hash_kernelILi6ENS_4Lay7EjE"), ("k1_flop_turn_keys", r"eval_stream_hash_kernelILi6ENS_5Lay56EjE"),
           ("k1_river_class16", r"eval_stream_hash_kernelILi6ENS_4Lay7EtE"), ("unrank_colex32", r"unrank_colex32_kernel"),
           ("k2_enum7_thread_walk", r"enum7_hash_kernel"), ("k2_enum7_warp_walk", r"enum7_warp_kernel"),
           ("k3_exhaustive_headsup", r"equity_exh_hash_kernelILi5ELi2ELb0"), ("k4_mc_9seats", r"equity_mc_fy_kernelILi8ELi768ELi1ELb0"),
           ("k4_mc_9seats_ev", r"equity_mc_fy_kernelILi8ELi768ELi1ELb1"), ("k5_preflop", r"preflop_hash_kernel"),
           ("k7b_board_count", r"pf_board_count_kernel"), ("mc_ranges", r"equity_mc_ranges_kernel")]
KEEP = re.compile(r"UBLKCP|SYNCS|LDG|STG|LDS|STS|ATOMS|ATOMG|REDG|RED\.|IMAD\.WIDE|IMAD\.HI|POPC|VIMNMX3|REDUX|BAR\.|LDC|WARPSYNC|SHFL|MATCH|VOTE")
names = subprocess.run(["cuobjdump", "-elf", SO], capture_output=True, text=True).stdout
syms = sorted(set(re.findall(r"(_ZN3mcp\w+)", names)))
for label, pat in KERNELS:
    fn = [s for s in syms if re.search(pat, s) and not s.endswith("_param")]
    if not fn:
        print("no symbol for", label)
        continue
    sass = subprocess.run(["cuobjdump", "-sass", "-fun", fn[0], SO], capture_output=True, text=True).stdout
    ins = [re.sub(r"\s*/\*[0-9a-fx]+\*/\s*$", "", l.split("*/", 1)[1]).strip() for l in sass.splitlines() if re.match(r"\s+/\*[0-9a-f]{4}\*/", l)]
    hist = collections.Counter(re.sub(r"^@!?U?P\d\s+", "", i).split()[0].rstrip(";") for i in ins if i)
    with open(os.path.join(out_dir, "%s_sass_%s.txt" % (tag, label)), "w") as f:
        f.write("# %s\n# cuobjdump -sass -fun %s monte-carlo_b200/libmcpoker.so  (sm_100a, %d instructions)\n" % (label, fn[0], len(ins)))
        f.write("# opcode histogram (static):\n")
        for op, c in hist.most_common():
            f.write("#   %-28s %d\n" % (op, c))
        f.write("# memory / TMA / barrier / wide-multiply instructions in program order:\n")
        for i in ins:
            if KEEP.search(i):
                f.write(i + "\n")
    print(label, len(ins), "instructions;", ", ".join("%s %d" % kv for kv in hist.most_common(6)))
