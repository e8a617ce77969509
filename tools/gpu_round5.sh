#!/bin/bash
# One gpurun call of round r05 (final state of this round): GPU parity tests, smoke, both bench arms, the five
# configs, brute-force variants and database-size sweep, launch list of the bench command, full captures of the
# extraction kernels and of the tcgen05 brute-force kernel.  Outputs: gpurun_out/<tag>_*.
TAG=${1:-r05}
mkdir -p gpurun_out
set -o pipefail
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${TAG}_pytest.log
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"; cat gpurun_out/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
echo "bench ref rc=$?"; cat gpurun_out/${TAG}_bench_ref.json
timeout 300 python tools/hamming_variants.py > gpurun_out/${TAG}_variants.txt 2>&1; cat gpurun_out/${TAG}_variants.txt
timeout 300 python tools/bf_sweep.py > gpurun_out/${TAG}_bf_sweep.txt 2>&1; cat gpurun_out/${TAG}_bf_sweep.txt
timeout 900 python tools/config_bench.py > gpurun_out/${TAG}_configs.log 2>&1; echo "configs rc=$?"; cp gpurun_out/configs.json gpurun_out/${TAG}_configs.json 2>/dev/null
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --batch 128 --skip-cpu --db-rows 2000000 > gpurun_out/${TAG}_ncu_launch.log 2>&1
echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'bruteforce_tc5' --launch-skip 2 -c 1 -f -o gpurun_out/${TAG}_ham \
    python tools/tc5_time.py 10000000 3 > gpurun_out/${TAG}_ham_ncu.log 2>&1
echo "ncu ham rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'^(blur|fast_cells|orient_desc|quadtree|resize_level)' --launch-skip 40 -c 22 -f -o gpurun_out/${TAG}_full \
    python bench.py --steps 2 --warmup 3 --batch 128 --skip-hamming --skip-cpu > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
python tools/shape_stages.py > gpurun_out/${TAG}_shape_stages.txt 2>&1; cat gpurun_out/${TAG}_shape_stages.txt
