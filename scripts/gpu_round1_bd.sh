#!/bin/bash
# lane-pair kernels compiled for more resident CTAs (forward 8, backward 6): c2, c3, suite
mkdir -p gpurun_out
B="timeout 300 python bench.py --no-cpu-baseline"
$B --workload c2 > gpurun_out/bd_c2.json 2> gpurun_out/bd.err
$B --workload c3 --steps 5 --warmup 3 > gpurun_out/bd_c3.json 2>> gpurun_out/bd.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bd_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -1
