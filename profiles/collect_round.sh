#!/bin/bash
# Collects the round's single-GPU evidence in one go (run on the GPU box from the repo root):
#   bash profiles/collect_round.sh r1   ->  gpurun_out/r1_*  (copy what should be judged into profiles/)
# Every ncu pass runs AFTER the same command exited 0 without ncu; numbers printed under ncu are never bench values.
set -u
tag=${1:-r1}
out=gpurun_out
mkdir -p $out
python bench.py > $out/${tag}_bench_base.json 2> $out/${tag}_bench_base.err || exit 1
python bench.py --variant grad --no-extra > $out/${tag}_bench_grad.json 2> $out/${tag}_bench_grad.err || exit 1
python bench.py --impl reference --steps 40 --warmup 3 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err
for v in base grad; do
  python profiles/track_phases.py $v > $out/${tag}_track_phases_$v.txt 2>&1
  python profiles/frame_log.py $v dev > $out/${tag}_frame_log_$v.txt 2>&1
  python profiles/kernel_times.py $v > $out/${tag}_warm_kernel_times_$v.txt 2>&1
  python profiles/sweep.py $v > $out/${tag}_sweep_$v.md 2> $out/${tag}_sweep_$v.err
done
# launch list of a short bench run (the same command just ran clean above with more steps)
python bench.py --steps 12 --warmup 3 --no-extra --cpu-budget-s 0.5 > $out/${tag}_bench_steps12.json 2>/dev/null &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches_bench_steps12.csv \
    python bench.py --steps 12 --warmup 3 --no-extra --cpu-budget-s 0.5 > $out/${tag}_ncu_launches.log 2>&1
python profiles/launch_summary.py $out/${tag}_launches_bench_steps12.csv > $out/${tag}_launch_summary.txt 2>&1
echo done
