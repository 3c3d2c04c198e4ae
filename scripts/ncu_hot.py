"""Top SASS instructions of a kernel by warp-stall samples, from an ncu report's source page.
  python scripts/ncu_hot.py gpurun_out/X.ncu-rep <kernel regex> [N]"""
import csv
import io
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
lines = txt.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
tot = sum(int(r["# Samples"] or 0) for r in rows) or 1
print(lines[0][:160])
print("total samples", tot, "instructions", len(rows))
idx = {id(r): i for i, r in enumerate(rows)}
for r in sorted(rows, key=lambda r: -int(r["# Samples"] or 0))[:n]:
    print("%5d %5.1f%%  #%-5d %s" % (int(r["# Samples"]), 100.0 * int(r["# Samples"]) / tot, idx[id(r)], r["Source"].strip()[:110]))
