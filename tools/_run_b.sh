set -o pipefail
P='import json,sys; d=json.loads(sys.stdin.read()); print("dev %.0f f/s %.3f ms | e2e %.0f f/s %.3f ms" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"]))'
run() { echo "--- $*"; env "$@" python bench.py --steps 20 --warmup 3 --skip-cpu --skip-hamming | python -c "$P"; }
run ORB_GPU_FORK_AT=2
run ORB_GPU_FORK_AT=1
run ORB_GPU_NO_FORK=1
run ORB_GPU_DEVICE_SPLIT=2
run ORB_GPU_DEVICE_SPLIT=4
run ORB_GPU_DEVICE_SPLIT=2 ORB_GPU_NO_FORK=1
run ORB_GPU_DEVICE_SPLIT=4 ORB_GPU_NO_FORK=1
run ORB_GPU_DEVICE_SPLIT=8 ORB_GPU_NO_FORK=1
python -m pytest tests/test_gpu_extractor.py tests/test_gpu_fullsize.py -m gpu -q -x 2>&1 | tail -3
ORB_GPU_DEVICE_SPLIT=4 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -x 2>&1 | tail -3
