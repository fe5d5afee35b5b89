#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2c}
{
for p in 0 100 200 400 800; do echo "GSDF_POLL_NS=$p"; GSDF_POLL_NS=$p timeout 200 python profiles/quick_fps.py grad; done
echo "GSDF_SERIAL=1"; GSDF_SERIAL=1 timeout 200 python profiles/quick_fps.py grad
echo "GSDF_BATCH_FRAMES=1"; GSDF_BATCH_FRAMES=1 timeout 200 python profiles/quick_fps.py grad
echo base; timeout 200 python profiles/quick_fps.py base
echo "base GSDF_POLL_NS=0"; GSDF_POLL_NS=0 timeout 200 python profiles/quick_fps.py base
} > $out/${tag}_ab.txt 2>&1
cat $out/${tag}_ab.txt
GSDF_POLL_NS=200 timeout 300 python profiles/frame_log.py grad dev > $out/${tag}_frame_log_grad.txt 2>&1; tail -n 1 $out/${tag}_frame_log_grad.txt
