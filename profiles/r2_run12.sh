#!/bin/bash
# 2 GPUs: sharding checks, then the bench line at N = 2
out=gpurun_out; mkdir -p $out; tag=${1:-r2l}; N=${2:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tests/multigpu_check.py > $out/${tag}_multigpu_check_n$N.log 2>&1; tail -n 2 $out/${tag}_multigpu_check_n$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $N --reps 3 > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo rc=$?
python - <<PY
import json
d=json.loads(open("$out/${tag}_bench_n$N.json").read().strip().splitlines()[-1]); print(d["value"], d["e2e"]["value"], d["shards_used"], d["sharded"]["value"]); r=d["room_scale"]; print({k:r[k] for k in ("frames_per_s","voxel_updates_per_s","converged","passes_per_frame","voxels")}, r["roofline"]["frac"])
PY
