#!/bin/bash
# Final evidence of a round in one gpurun call: tools/gpu_round.sh (tests, smoke, both bench arms, launch list), then
# the full ncu captures (extraction step at batch 256, brute force on 2 M rows), the secondary configs, the
# brute-force size sweep and variants, the blur A/B.  usage: tools/gpu_final.sh <tag>
TAG=${1:-r10}
bash tools/gpu_round.sh $TAG
SKIP=44 COUNT=11 bash tools/gpu_ncu.sh $TAG 256 '^(blur|fast_cells|orient_desc|quadtree|resize_level)'
bash tools/gpu_ncu_hamming.sh $TAG 2000000
timeout 600 python tools/config_bench.py > gpurun_out/${TAG}_configs.log 2>&1; echo "configs rc=$?"; cp gpurun_out/configs.json gpurun_out/${TAG}_configs.json 2>/dev/null
timeout 300 python tools/bf_sweep.py 10000000 5000000 2500000 1250000 625000 100000 > gpurun_out/${TAG}_bf_sweep.txt 2>&1; echo "sweep rc=$?"
timeout 300 python tools/hamming_variants.py > gpurun_out/${TAG}_variants.txt 2>&1; echo "variants rc=$?"
( echo "== default (tensor-core blur)"; timeout 100 python tools/blur_check.py tum 2>&1 | tail -4
  echo "== ORB_GPU_BLUR=simd"; ORB_GPU_BLUR=simd timeout 100 python tools/blur_check.py tum 2>&1 | tail -4 ) > gpurun_out/${TAG}_blur_ab.txt 2>&1
timeout 200 python tools/shape_stages.py > gpurun_out/${TAG}_shape_stages.txt 2>&1; echo "shapes rc=$?"
