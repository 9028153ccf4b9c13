#!/bin/bash
# round-2 ncu evidence on one B200 (run under gpurun; every command first exits 0 without ncu):
#   1. launch list (gpu__time_duration per launch) of this library's kernels inside the bench command
#   2. --set full of the band solver's kernels and of the global-BA build at configs[3] size
O=gpurun_out/prof_r2
mkdir -p $O
B="python bench.py --no-cpu --no-ref-cuda --no-sharded --quick --steps 10 --warmup 30 --breakdown-frames 0"
$B > $O/bench_plain.json 2> $O/bench_plain.err || exit 1
# this library's kernels inside the bench command (-k filters by base name, everything else runs unprofiled: an all-kernel pass
# took > 30 min, the clip renderer alone launches ~50 k elementwise kernels); -c 1700 = the 40 tracked frames (43.6 launches per
# frame, bench line `gpu_launches_per_frame`), before the hot-path micro-benchmarks of the same command start
OURS='regex:corr_|restage|ba_window|ba_reproject|plan_|lie_|pops_|mirror_words|neighbors|soft_agg|group_plan|frame_distance|patchify'
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" -c 1700 --csv --log-file $O/launches_bench_all.csv $B > $O/bench_under_ncu.json 2> $O/bench_under_ncu.err
if [ -z "$SKIP_BAND" ]; then
python tools/band_one.py > $O/band_one.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'ba_band|ba_global_kernel|ba_global_span' -c 8 -o $O/band_global_r2 python tools/band_one.py > $O/ncu_band.log 2>&1
fi
python - <<'PY'
import csv, collections, json
rows = list(csv.reader(l for l in open("gpurun_out/prof_r2/launches_bench_all.csv") if l.startswith('"')))
h = rows[0]; k = h.index("Kernel Name"); v = h.index("Metric Value")
tot = collections.OrderedDict()
for r in rows[1:]:
    name = r[k].split("(")[0].replace("wv::", "").replace("void ", "")
    t = tot.setdefault(name, [0, 0.0]); t[0] += 1; t[1] += float(r[v].replace(",", "")) / 1e3
allus = sum(t[1] for t in tot.values())
d = json.loads(open("gpurun_out/prof_r2/bench_plain.json").read().strip().splitlines()[-1])
frames = d["steps"] + d["warmup"]
with open("gpurun_out/prof_r2/launches_bench_summary.txt", "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -k <kernels of libwvb200.so> -c 1700 over\n")
    f.write("#   python bench.py --no-cpu --no-ref-cuda --no-sharded --quick --steps 10 --warmup 30 --breakdown-frames 0\n")
    f.write(f"# {len(rows) - 1} launches of this library's kernels over the {frames} tracked frames: {allus / 1e3:.2f} ms in total = {allus / 1e3 / frames:.3f} ms per frame\n")
    f.write(f"# the same command without ncu: {d['value']:.1f} frames/s = {d['ms_per_step']:.2f} ms per frame (CUDA events) -> this library's kernels are {100 * allus / 1e3 / frames / d['ms_per_step']:.1f} % of a frame;\n")
    f.write("# the rest is the network (cuDNN / cuBLAS / elementwise), the harness's torch_scatter shim and host code (bench line `ms_breakdown`)\n")
    f.write("kernel, launches, launches per frame, total us, avg us, share of this library's kernel time\n")
    for n, (c, us) in sorted(tot.items(), key=lambda x: -x[1][1]):
        f.write(f"{n}, {c}, {c / frames:.1f}, {us:.1f}, {us / c:.2f}, {100 * us / allus:.1f}%\n")
print(open("gpurun_out/prof_r2/launches_bench_summary.txt").read()[:3500])
PY
rm -f $O/launches_bench_all.csv
