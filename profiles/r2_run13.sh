#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2m}
timeout 900 python -m pytest tests/test_host_cpp.py -m gpu -x -q > $out/${tag}_pytest_host.log 2>&1; echo "rc=$?" >> $out/${tag}_pytest_host.log
tail -n 25 $out/${tag}_pytest_host.log
