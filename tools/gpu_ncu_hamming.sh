#!/bin/bash
# ncu full capture of the brute-force matcher kernels.  usage: tools/gpu_ncu_hamming.sh <tag> [db rows]
TAG=${1:-ham}; N=${2:-2000000}
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'bruteforce_tc5' --launch-skip 2 -c 2 -f -o gpurun_out/${TAG}_ham \
    python tools/hamming_variants.py $N > gpurun_out/${TAG}_ham_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/${TAG}_ham_ncu.log
