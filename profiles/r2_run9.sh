#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2i}
timeout 1200 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "rc=$?" >> $out/${tag}_pytest.log
tail -n 4 $out/${tag}_pytest.log
GSDF_DBG_SKIP=50 timeout 300 python profiles/track_phases.py grad > $out/${tag}_track_phases_grad.txt 2>&1; tail -n 32 $out/${tag}_track_phases_grad.txt
