#!/bin/bash
# One gpurun call: GPU parity tests, bench, ncu launch list, ncu full capture.  usage: tools/gpu_round.sh <tag> [batch]
# Outputs go to gpurun_out/<tag>_*.  Numbers printed under ncu are never bench values.
TAG=${1:-r01}
B=${2:-512}
mkdir -p gpurun_out
set -o pipefail
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${TAG}_pytest.log
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 --batch $B > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"; cat gpurun_out/${TAG}_bench.json
if [ "$3" != "noncu" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --batch 128 --skip-hamming --skip-cpu > gpurun_out/${TAG}_ncu_launch.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'^(blur|fast_cells|orient_desc|quadtree|resize_level)' --launch-skip 40 -c 20 -f -o gpurun_out/${TAG}_full \
    python bench.py --steps 2 --warmup 3 --batch 128 --skip-hamming --skip-cpu > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
fi
