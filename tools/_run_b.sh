set -o pipefail
timeout 600 python -m pytest tests/test_gpu_matcher.py tests/test_gpu_fullsize.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -15
timeout 300 python tools/bf_sweep.py 10000000 1250000 2>&1 | tail -2
