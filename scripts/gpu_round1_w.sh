#!/bin/bash
# group-major gradient hand-off head -> hash-grid backward: suite + c5 A/B
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B --workload c5 > gpurun_out/w_c5.json 2>> gpurun_out/w.err
$B --workload c5 --no-grouped-head > gpurun_out/w_c5_rows.json 2>> gpurun_out/w.err
$B --workload c3 > gpurun_out/w_c3.json 2>> gpurun_out/w.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/w_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['config'].get('head_grad_group_width'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/w.err
