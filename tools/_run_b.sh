set -o pipefail
echo "== alone"; python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -3
echo "== all"; python -m pytest tests -m gpu -q -x 2>&1 | tail -4
python bench.py --steps 10 --warmup 3 --skip-hamming --skip-cpu --skip-parity 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('latency', d['latency_us_batch1'], 'value', d['value'], 'e2e', d['e2e']['value'], d['roofline']['stage_ms'])"
