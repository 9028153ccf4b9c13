"""Per-source-line warp-stall sample shares from an ncu report (needs -lineinfo and --import-source on).

    python tools/ncu_lines.py report.ncu-rep [top_n] [launch_index]
"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
cur, hdr, agg, k = None, None, {}, None
for r in csv.reader(txt.splitlines()):
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) == 2 and r[0] == "Function Name":
        print("#", r[1])
    elif len(r) > 4 and r[0] == "Line No":
        hdr = r
        k = hdr.index("Warp Stall Sampling (All Samples)")
    elif hdr and len(r) == len(hdr) and r[0] != "":
        try:
            agg[(cur, int(r[0]))] = (float(r[k]), r[1])
        except ValueError:
            pass
tot = sum(v for v, _ in agg.values()) or 1.0
print("# total samples", tot)
for (f, l), (v, src) in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print("%5.1f%% %s:%d  %s" % (100 * v / tot, f, l, src.strip()[:110]))
