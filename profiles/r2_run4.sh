#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2d}
timeout 600 python -m pytest tests/test_outputs.py tests/test_host_cpp.py -m gpu -x -q > $out/${tag}_pytest_outputs.log 2>&1; echo "rc=$?" >> $out/${tag}_pytest_outputs.log
tail -n 30 $out/${tag}_pytest_outputs.log
