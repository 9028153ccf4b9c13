#!/bin/bash
# the multi-GPU measurements of round 2 on N GPUs of one box: sharded global BA (band and dense routes, with the 1-GPU comparison),
# bench.py's own line, the clip batch.  Usage: tools/multi_gpu_r2.sh N   (run under gpurun --gpus N)
N=${1:-2}
O=gpurun_out/mgpu_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 tools/global_ba.py --bench > $O/global_ba_bench.json 2> $O/global_ba_bench.err
$TR --master-port 29512 tools/global_ba.py --frames 2048 --M 192 --edges config4 --check > $O/global_ba_check_config4.json 2> $O/global_ba_check.err
$TR --master-port 29513 tools/global_ba.py --frames 2048 --M 96 --edges global --check > $O/global_ba_check_global.json 2>> $O/global_ba_check.err
WVB200_GLOBAL_DENSE=1 $TR --master-port 29514 tools/global_ba.py --frames 2048 --M 192 --edges config4 --check > $O/global_ba_check_config4_dense.json 2>> $O/global_ba_check.err
$TR --master-port 29515 bench.py --gpus $N --no-cpu > $O/bench.json 2> $O/bench.err
tail -c 400 $O/global_ba_bench.json; echo; cat $O/global_ba_check_config4.json | cut -c1-900; echo; cat $O/global_ba_check_global.json | cut -c1-900; echo;  cat $O/global_ba_check_config4_dense.json | cut -c1-900; echo; python - <<PY
import json
d = json.loads(open("$O/bench.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("metric", "value", "n_gpus", "ms_per_step", "scaling")}, d.get("e2e"), d.get("sharded", {}).get("ms_2_iterations"), d.get("ref_cuda", {}).get("value"))
PY
tail -n 3 $O/bench.err | cut -c1-300
# configs[4]: a clip batch over the N GPUs (8 clips per GPU), then one clip of every rank's shard again in ONE process: the COLMAP
# export must not depend on the sharding (trees stay on the box; the JSON lines come back)
C=$((8 * N)); FR=${CLIP_FRAMES:-80}
$TR --master-port 29516 tools/run_clips.py --clips $C --frames $FR --out /tmp/clips_n$N > $O/clips.json 2> $O/clips.err
ONLY=$(python -c "print(','.join(str(8 * r + (r % 8)) for r in range($N)))")
python tools/run_clips.py --only $ONLY --frames $FR --out /tmp/clips_one > $O/clips_one_process.json 2>> $O/clips.err
python tools/run_clips.py --compare /tmp/clips_n$N /tmp/clips_one --subset > $O/clips_compare.json 2>> $O/clips.err
python - <<PY
import json
d = json.loads(open("$O/clips.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("metric", "value", "n_gpus", "clips", "frames_per_clip", "seconds", "ate_rmse_m")})
print(open("$O/clips_compare.json").read())
PY
tail -n 3 $O/clips.err | cut -c1-300
