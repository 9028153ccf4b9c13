"""Per-SASS-instruction warp-stall samples with their reasons from an ncu report (--import-source on).

    python tools/ncu_sass.py report.ncu-rep [top_n] [launch_index]
"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
print("#", rows[0][1])
h = rows[1]
ix = {k: i for i, k in enumerate(h)}
R = [r for r in rows[2:] if len(r) == len(h) and r[0].startswith("0x")]
S = "Warp Stall Sampling (All Samples)"
f = lambda r, k: float(r[ix[k]] or 0)  # noqa: E731
tot = sum(f(r, S) for r in R) or 1.0
print("# instructions", len(R), "samples", tot)
keys = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
top = sorted(range(len(R)), key=lambda i: -f(R[i], S))[:top_n]
for i in sorted(top):
    r = R[i]
    why = " ".join(f"{k[6:]}={int(f(r, k))}" for k in keys if f(r, k))
    print(f"{i:5d} {r[ix['Source']].strip()[:64]:64s} {f(r, S):5.0f} x{int(f(r, 'Instructions Executed')):7d}  {why}")
