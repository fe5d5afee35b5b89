#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2b}
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest.log
tail -n 5 $out/${tag}_pytest.log
for v in grad base; do
  timeout 300 python profiles/track_phases.py $v > $out/${tag}_track_phases_${v}.txt 2>&1
  timeout 300 python profiles/frame_log.py $v dev > $out/${tag}_frame_log_${v}.txt 2>&1
  tail -n 1 $out/${tag}_frame_log_${v}.txt
done
timeout 600 python bench.py --variant grad --no-extra --cpu-budget-s 1 > $out/${tag}_bench_grad.json 2> $out/${tag}_bench_grad.err; tail -c 600 $out/${tag}_bench_grad.json
