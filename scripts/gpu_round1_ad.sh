#!/bin/bash
# 256-bit loads / stores in the head kernels: suite + c5
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B --workload c5 > gpurun_out/ae_c5.json 2>> gpurun_out/ae.err
$B --workload c5 --no-grouped-head > gpurun_out/ae_c5_rows.json 2>> gpurun_out/ae.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/ad_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/ae.err
