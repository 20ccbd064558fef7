#!/bin/bash
# nearest-cell default of 8 levels per group: parity + c1r bench line (with cpu baseline) for profiles/
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/al_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/al_pytest.log
tail -3 gpurun_out/al_pytest.log
timeout 400 python bench.py --workload c1r > gpurun_out/al_c1r.json 2>> gpurun_out/al.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/al_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['roofline_packets']['forward']['frac'], d['roofline_packets']['backward']['frac'])
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/al.err
