#!/bin/bash
# nearest-cell (Mode R) kernels with two points per trip: parity + c1r bench
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/ai_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ai_pytest.log
tail -3 gpurun_out/ai_pytest.log
B="timeout 400 python bench.py --no-cpu-baseline"
$B --workload c1r > gpurun_out/ai_c1r.json 2>> gpurun_out/ai.err
$B --workload c1r > gpurun_out/ai_c1r_b.json 2>> gpurun_out/ai.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/ai_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/ai.err
