#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2g}
for v in grad base; do
  timeout 300 python profiles/frame_log.py $v dev > $out/${tag}_frame_log_${v}.txt 2>&1; tail -n 2 $out/${tag}_frame_log_${v}.txt
  GSDF_SERIAL=1 timeout 300 python profiles/frame_log.py $v dev > $out/${tag}_frame_log_${v}_serial.txt 2>&1; tail -n 2 $out/${tag}_frame_log_${v}_serial.txt
done
timeout 200 python profiles/quick_fps.py grad; timeout 200 python profiles/quick_fps.py base
