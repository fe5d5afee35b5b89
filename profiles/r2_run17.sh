#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2s}
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "rc=$?" >> $out/${tag}_pytest.log
tail -n 6 $out/${tag}_pytest.log
for v in grad base; do timeout 300 python profiles/frame_log.py $v dev > $out/${tag}_frame_log_$v.txt 2>&1; tail -n 2 $out/${tag}_frame_log_$v.txt; done
GSDF_SERIAL=1 timeout 300 python profiles/frame_log.py grad dev | tail -n 2
timeout 200 python profiles/quick_fps.py grad; timeout 200 python profiles/quick_fps.py base
timeout 300 python profiles/track_phases.py grad steady > $out/${tag}_track_phases_grad.txt 2>&1; sed -n 2,8p $out/${tag}_track_phases_grad.txt; grep -A6 "steady state" $out/${tag}_track_phases_grad.txt | head -8
