#!/bin/bash
# Round 2 evidence, the commands behind the files under profiles/ (run on a GPU box from the repo root, e.g.
#   gpurun --timeout 3000 -- 'bash profiles/collect_round2.sh r2z').  Every ncu pass follows a plain run of the same command
# that exited 0; numbers printed under ncu are never bench values.
set -u
tag=${1:-r2z}; out=gpurun_out; mkdir -p $out
# 1. parity tests, both bench arms, phase stamps, steady-state frame anatomy            (= profiles/r2_final1.sh)
bash profiles/r2_final1.sh $tag
# 2. ncu: --set full captures of the tracker (grad, base) and the frame-only kernel with the reduction counters, launch list
bash profiles/r2_ncu.sh ${tag}n
# 3. A/B switches: shared-memory carveout, polling pauses / serial mode / one frame per launch
for c in 100 64 50 40 32; do echo "GSDF_CARVEOUT=$c"; GSDF_CARVEOUT=$c python profiles/quick_fps.py grad; done > $out/${tag}_carveout_sweep.txt 2>&1
{ for p in 0 200 800; do echo "GSDF_POLL_NS=$p"; GSDF_POLL_NS=$p python profiles/quick_fps.py grad; done
  echo "GSDF_SERIAL=1"; GSDF_SERIAL=1 python profiles/quick_fps.py grad
  echo "GSDF_BATCH_FRAMES=1"; GSDF_BATCH_FRAMES=1 python profiles/quick_fps.py grad; } > $out/${tag}_ab.txt 2>&1
# 4. run-to-run behaviour: tracker bit-reproducible on a fixed map; spread of the noisy C1 loop
python profiles/r2_determinism_check.py > $out/${tag}_determinism.txt 2>&1
python profiles/r2_c1_repeat.py 4 > $out/${tag}_c1_repeat.txt 2>&1
# N > 1 (gpurun --gpus N -- 'bash profiles/r2_scale.sh <tag> N'): sharding checks + bench line at N GPUs
echo done
