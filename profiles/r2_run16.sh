#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2r}
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "rc=$?" >> $out/${tag}_pytest.log
tail -n 30 $out/${tag}_pytest.log
timeout 200 python profiles/quick_fps.py grad; timeout 200 python profiles/quick_fps.py base
