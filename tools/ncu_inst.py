"""Warp-instruction counts per source line and per source region from an ncu report (needs -lineinfo and --import-source on).

    python tools/ncu_inst.py report.ncu-rep launch_index units [file:lo-hi=name ...]
units: the number of work items of the launch (counts are printed per item)."""
import csv, subprocess, sys
rep, skip, units = sys.argv[1], sys.argv[2], float(sys.argv[3])
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
cur, hdr, agg, wav = None, None, {}, {}
for r in csv.reader(txt.splitlines()):
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 4 and r[0] == "Line No":
        hdr = r
        ki, kw = hdr.index("Instructions Executed"), hdr.index("L1 Wavefronts Shared")
    elif hdr and len(r) == len(hdr) and r[0] != "":
        try:
            agg[(cur, int(r[0]))] = (float(r[ki]), r[1]); wav[(cur, int(r[0]))] = float(r[kw] or 0)
        except ValueError:
            pass
tot = sum(v for v, _ in agg.values())
print(f"total warp instructions {tot:.0f} = {tot / units:.1f} per item; shared wavefronts {sum(wav.values()) / units:.1f} per item")
for spec in sys.argv[4:]:
    loc, name = spec.split("=")
    f, rng = loc.split(":"); a, b = map(int, rng.split("-"))
    s = sum(v for (ff, l), (v, _) in agg.items() if ff == f and a <= l <= b)
    w = sum(v for (ff, l), v in wav.items() if ff == f and a <= l <= b)
    print(f"  {name:28s} {s / units:8.1f} instr  {w / units:8.1f} smem wavefronts")
for (f, l), (v, src) in sorted(agg.items(), key=lambda x: -x[1][0])[:30]:
    print(f"{v / units:7.1f} {wav[(f, l)] / units:7.1f}  {f}:{l} {src[:100]}")
