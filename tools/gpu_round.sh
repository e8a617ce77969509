#!/bin/bash
# One gpurun call: GPU parity tests, smoke, both bench arms, launch list of the bench command.
# Outputs: gpurun_out/<tag>_*.   Usage: tools/gpu_round.sh <tag> [pytest-args...]
TAG=${1:-r2}
shift
mkdir -p gpurun_out
set -o pipefail
timeout 1500 python -m pytest tests -m gpu -x -q "$@" 2>&1 | tail -25 > gpurun_out/${TAG}_pytest.log
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -8 gpurun_out/${TAG}_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"; cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
if [ -z "$SKIP_REF" ]; then
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err
echo "ref rc=$?"; cat gpurun_out/${TAG}_bench_ref.json
fi
if [ -z "$SKIP_NCU" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --batch 128 --skip-cpu --db-rows 2000000 > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
fi
