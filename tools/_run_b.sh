set -o pipefail
python -m pytest tests/test_gpu_matcher.py tests/test_gpu_bow_init.py tests/test_gpu_orbmatcher_class.py tests/test_gpu_vs_reference.py -m gpu -q -x 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --skip-hamming --skip-cpu --skip-parity 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(json.dumps(d['configs']['search_by_projection'],indent=1)); print(json.dumps(d['configs']['search_by_bow'],indent=1))"
