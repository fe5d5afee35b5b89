#!/bin/bash
# one N of the scaling series: sharding checks (N >= 2) and the bench line
out=gpurun_out; mkdir -p $out; tag=${1:-r2z}; N=${2:-2}
if [ "$N" -ge 2 ]; then
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 tests/multigpu_check.py > $out/${tag}_multigpu_check_n$N.log 2>&1; tail -n 1 $out/${tag}_multigpu_check_n$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29553 bench.py --gpus $N --reps 5 > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo rc=$?
else
python bench.py --reps 5 > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo rc=$?
fi
python - <<PY
import json
d=json.loads(open("$out/${tag}_bench_n$N.json").read().strip().splitlines()[-1]); print("N=$N value",d["value"],"e2e",d["e2e"]["value"],"shards_used",d.get("shards_used"),"sharded",(d.get("sharded") or {}).get("value")); r=d["room_scale"]; print(" room",r["frames_per_s"],r["voxel_updates_per_s"],r["converged"],r["roofline"]["frac"])
PY
