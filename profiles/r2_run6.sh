#!/bin/bash
out=gpurun_out; mkdir -p $out; tag=${1:-r2f}
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "noisy_grad or reference_build_live" > $out/${tag}_pytest_new.log 2>&1; echo "rc=$?" >> $out/${tag}_pytest_new.log
tail -n 12 $out/${tag}_pytest_new.log
timeout 900 python bench.py > $out/${tag}_bench_grad.json 2> $out/${tag}_bench_grad.err; echo "bench rc=$?"; tail -c 1500 $out/${tag}_bench_grad.json; tail -n 5 $out/${tag}_bench_grad.err
timeout 600 python bench.py --impl reference --steps 40 --warmup 3 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; tail -c 700 $out/${tag}_bench_reference.json
