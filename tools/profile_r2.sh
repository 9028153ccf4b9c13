#!/bin/bash
# round-2 ncu evidence on one B200 (run under gpurun; every command first exits 0 without ncu):
#   1. launch list (gpu__time_duration per launch) of this library's kernels inside the bench command
#   2. --set full of the band solver's kernels and of the global-BA build at configs[3] size
O=gpurun_out/prof_r2
mkdir -p $O
B="python bench.py --no-cpu --no-ref-cuda --no-sharded --quick --steps 20 --warmup 30 --breakdown-frames 0"
$B > $O/bench_plain.json 2> $O/bench_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'wv::' --csv --log-file $O/launches_bench.csv $B > $O/bench_under_ncu.json 2> $O/bench_under_ncu.err
python tools/band_one.py > $O/band_one.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'ba_band|ba_global_kernel|ba_global_span' -c 8 -o $O/band_global_r2 python tools/band_one.py > $O/ncu_band.log 2>&1
python - <<'PY'
import csv, collections, json
rows = list(csv.reader(l for l in open("gpurun_out/prof_r2/launches_bench.csv") if l.startswith('"')))
h = rows[0]; k = h.index("Kernel Name"); v = h.index("Metric Value")
tot = collections.OrderedDict()
for r in rows[1:]:
    name = r[k].split("(")[0]
    t = tot.setdefault(name, [0, 0.0]); t[0] += 1; t[1] += float(r[v].replace(",", "")) / 1e3
allus = sum(t[1] for t in tot.values())
print("kernel, launches, total us, avg us, share")
for n, (c, us) in sorted(tot.items(), key=lambda x: -x[1][1]):
    print(f"{n}, {c}, {us:.1f}, {us / c:.2f}, {100 * us / allus:.1f}%")
d = json.loads(open("gpurun_out/prof_r2/bench_plain.json").read().strip().splitlines()[-1])
print("bench (no ncu):", d["value"], "frames/s; hot path", d["hotpath"])
PY
