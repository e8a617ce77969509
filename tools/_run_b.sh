set -o pipefail
python -m pytest tests/test_gpu_extractor.py tests/test_gpu_dropin.py tests/test_gpu_stereo.py -m gpu -q -x 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --skip-hamming --skip-cpu --skip-parity 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('latency', d['latency_us_batch1'], 'value', d['value'], 'e2e', d['e2e']['value']); print(json.dumps(d['configs']['kitti_stereo_pair'])[:300])"
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/r2d_smoke_launches.csv python __graft_entry__.py --smoke > /dev/null 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r2d_smoke_launches.csv')) if len(r)>5]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
for r in rows[1:16]:
    print(r[ki][:50].ljust(52), r[vi])
PY
