#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2k}
timeout 900 python bench.py --reps 3 --no-extra --cpu-budget-s 1 > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "rc=$?"; tail -n 3 $out/${tag}_bench_n1.err
python - <<PY
import json
d=json.load(open("$out/${tag}_bench_n1.json")); print(d["value"], d["e2e"]["value"], d.get("shards_used")); print(json.dumps(d.get("room_scale"))[:900])
PY
