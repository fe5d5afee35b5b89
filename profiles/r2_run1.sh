#!/bin/bash
# round 2, first GPU pass: parity tests with the new tracker exchange, then phase stamps for both CTA shapes
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> $out/r2a_pytest.log
tail -5 $out/r2a_pytest.log
for t in 768 384; do for v in grad base; do
  GSDF_TRACK_THREADS=$t timeout 300 python profiles/track_phases.py $v > $out/r2a_track_phases_${v}_$t.txt 2>&1
  GSDF_TRACK_THREADS=$t timeout 300 python profiles/frame_log.py $v dev > $out/r2a_frame_log_${v}_$t.txt 2>&1
done; done
tail -4 $out/r2a_frame_log_grad_768.txt $out/r2a_frame_log_grad_384.txt
