"""Summarise an .ncu-rep (read here, without a GPU) into the few counters DESIGN.md argues from.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [units-per-launch by kernel substring: name=units ...]"""
import csv
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("sm__inst_executed.avg.per_cycle_active", "IPC per SM (max 4)"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU pipe % of peak"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA pipe % of peak"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU pipe % of peak (POPC/FLO)"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU pipe % of peak"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of peak"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__grid_size", "grid"),
    ("sm__cycles_elapsed.max", "SM cycles"),
]


def main():
    rep = sys.argv[1]
    units = dict(a.split("=") for a in sys.argv[2:])
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, unit_row = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    seen = set()
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        if name in seen:
            continue
        seen.add(name)
        print("== %s" % name)
        for key, label in WANT:
            if key in col:
                print("   %-34s %s %s" % (label, r[col[key]], unit_row[col[key]]))
        for sub, u in units.items():
            if sub in name:
                wi = float(r[col["smsp__inst_executed.sum"]])
                print("   %-34s %.1f  (%s units per launch)" % ("warp-instr x32 per unit", wi * 32 / float(u), u))


if __name__ == "__main__":
    main()
