"""Writes profiles/k1_traffic.json from an `ncu --set full` capture of the K1 launch at the bench size:
dram__bytes_read.sum + dram__bytes_write.sum of eval_stream_hash_kernel<6, Lay7, unsigned int>, tagged with the SHA-256 of
the K1 sources (bench.k1_source_hash) so that bench.py only reports the traffic of the build it is benching.
usage: python tools/k1_traffic.py gpurun_out/prof.ncu-rep [batch_log2=28]"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

rep = sys.argv[1]
log2 = int(sys.argv[2]) if len(sys.argv) > 2 else 28
# an .ncu-rep, or the CSV `ncu -i rep --page raw --csv` made of it on the GPU box (reports do not fit gpurun_out/)
raw = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}


def to_bytes(v, u):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]


best = None
for r in rows[2:]:
    name = r[col["Kernel Name"]]
    if "eval_stream_hash_kernel<6, " in name and "Lay7, unsigned int>" in name:
        rd = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
        wr = to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
        best = (name, rd, wr)                      # the last launch of the capture
if not best:
    raise SystemExit("no K1 (Lay7, uint32) launch in %s" % rep)
out = {"kernel": best[0], "source": "ncu --set full, %s (2^%d hands, one launch)" % (os.path.basename(rep), log2),
       "batch_log2": log2, "dram_bytes_read": best[1], "dram_bytes_write": best[2],
       "dram_bytes_per_launch_at_bench_size": best[1] + best[2], "algorithmic_bytes_per_launch": 12 * (1 << log2),
       "k1_source_sha256": bench.k1_source_hash()}
json.dump(out, open(os.path.join(ROOT, "profiles", "k1_traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
