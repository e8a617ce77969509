#!/bin/bash
# ncu full capture of one step of the extraction kernels.  usage: tools/gpu_ncu.sh <tag> [batch] [kernel-regex]
TAG=${1:-r01}; B=${2:-128}; K=${3:-'^(blur|fast_cells|orient_desc|quadtree|resize_level)'}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$K" --launch-skip ${SKIP:-54} -c ${COUNT:-18} -f -o gpurun_out/${TAG}_full \
    python bench.py --steps 2 --warmup 3 --batch $B --skip-hamming --skip-cpu > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/${TAG}_ncu_full.log | cut -c1-300
