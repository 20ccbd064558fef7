#!/bin/bash
# head kernels with all tile loads in flight; e2e loop with overlapped input copies
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B --workload c5 > gpurun_out/mm_c5.json 2>> gpurun_out/mm.err
$B --workload c2 > gpurun_out/mm_c2.json 2>> gpurun_out/mm.err
$B --workload c3 > gpurun_out/mm_c3.json 2>> gpurun_out/mm.err
$B --workload c1r > gpurun_out/mm_c1r.json 2>> gpurun_out/mm.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/mm_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/mm.err
