#!/bin/bash
# round-2 ncu evidence on one B200 (run under gpurun; every command first exits 0 without ncu):
#   1. launch list (gpu__time_duration per launch) of this library's kernels inside the bench command
#   2. --set full of the band solver's kernels and of the global-BA build at configs[3] size
O=gpurun_out/prof_r2
mkdir -p $O
B="python bench.py --no-cpu --no-ref-cuda --no-sharded --quick --steps 10 --warmup 30 --breakdown-frames 0"
$B > $O/bench_plain.json 2> $O/bench_plain.err || exit 1
# every kernel of the command (torch's, cuDNN's, cuBLAS's and this library's) over ~8 steady-state frames (the launches before
# them run unprofiled): the library's share of a frame is the point
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip ${NCU_SKIP:-40000} --launch-count ${NCU_COUNT:-12000} --csv --log-file $O/launches_bench_all.csv $B > $O/bench_under_ncu.json 2> $O/bench_under_ncu.err
if [ -z "$SKIP_BAND" ]; then
python tools/band_one.py > $O/band_one.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'ba_band|ba_global_kernel|ba_global_span' -c 8 -o $O/band_global_r2 python tools/band_one.py > $O/ncu_band.log 2>&1
fi
python - <<'PY'
import csv, collections, json
OURS = ("corr_", "restage", "ba_", "plan_", "lie_", "pops_", "soft_agg", "group_", "frame_distance", "mirror_words", "gridsync", "patchify_", "neighbors")
rows = list(csv.reader(l for l in open("gpurun_out/prof_r2/launches_bench_all.csv") if l.startswith('"')))
h = rows[0]; k = h.index("Kernel Name"); v = h.index("Metric Value")
tot = collections.OrderedDict()
for r in rows[1:]:
    name = r[k].split("(")[0].split("<")[0].replace("wv::", "")
    t = tot.setdefault(name, [0, 0.0]); t[0] += 1; t[1] += float(r[v].replace(",", "")) / 1e3
allus = sum(t[1] for t in tot.values())
ours = sum(t[1] for n, t in tot.items() if n.startswith(OURS))
with open("gpurun_out/prof_r2/launches_bench_summary.txt", "w") as f:
    f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none over `python bench.py --no-cpu --no-ref-cuda --no-sharded --quick --steps 10 --warmup 30 --breakdown-frames 0`\n")
    f.write(f"# {len(rows) - 1} launches, {allus / 1e3:.1f} ms of kernel time in total; kernels of libwvb200.so: {ours / 1e3:.1f} ms = {100 * ours / allus:.1f} %\n")
    f.write("kernel, launches, total us, avg us, share of all kernel time, ours\n")
    for n, (c, us) in sorted(tot.items(), key=lambda x: -x[1][1])[:60]:
        f.write(f"{n}, {c}, {us:.1f}, {us / c:.2f}, {100 * us / allus:.2f}%, {'*' if n.startswith(OURS) else ''}\n")
print(open("gpurun_out/prof_r2/launches_bench_summary.txt").read()[:3500])
d = json.loads(open("gpurun_out/prof_r2/bench_plain.json").read().strip().splitlines()[-1])
print("bench (no ncu):", d["value"], "frames/s; hot path", d["hotpath"])
PY
rm -f $O/launches_bench_all.csv
