#!/bin/bash
for c in 100 64 50 40 32; do echo "GSDF_CARVEOUT=$c"; GSDF_CARVEOUT=$c timeout 200 python profiles/quick_fps.py grad; GSDF_CARVEOUT=$c timeout 200 python profiles/frame_log.py grad dev | tail -n 2; done
echo "base 50"; GSDF_CARVEOUT=50 timeout 200 python profiles/quick_fps.py base
