#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2o}
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "rc=$?" >> $out/${tag}_pytest.log
tail -n 4 $out/${tag}_pytest.log
timeout 300 python profiles/frame_log.py grad dev > $out/${tag}_frame_log_grad.txt 2>&1; tail -n 2 $out/${tag}_frame_log_grad.txt
timeout 200 python profiles/quick_fps.py grad; timeout 200 python profiles/quick_fps.py base
