#!/bin/bash
# final single-GPU evidence of round 2
out=gpurun_out; mkdir -p $out; tag=${1:-r2z}
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "rc=$?" >> $out/${tag}_pytest_gpu.log; tail -n 3 $out/${tag}_pytest_gpu.log
timeout 900 python bench.py --impl reference --steps 40 --warmup 3 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; echo "ref rc=$?"
timeout 1200 python bench.py > $out/${tag}_bench_grad.json 2> $out/${tag}_bench_grad.err; echo "bench rc=$?"
for v in grad base; do
  timeout 300 python profiles/track_phases.py $v steady > $out/${tag}_track_phases_$v.txt 2>&1
  timeout 300 python profiles/frame_log.py $v dev > $out/${tag}_frame_log_$v.txt 2>&1; tail -n 2 $out/${tag}_frame_log_$v.txt
done
GSDF_SERIAL=1 timeout 300 python profiles/frame_log.py grad dev > $out/${tag}_frame_log_grad_serial.txt 2>&1; tail -n 2 $out/${tag}_frame_log_grad_serial.txt
python -c "
import json
d=json.load(open('$out/${tag}_bench_grad.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['cpu_baseline']['value'], d['room_scale']['frames_per_s'], d['room_scale']['roofline']['frac'], d['extra']['frames_per_s'])"
