#!/bin/bash
# Mode R (c1r): backward levels per group sweep, forward at 8
set -x
mkdir -p gpurun_out
B="timeout 400 python bench.py --no-cpu-baseline --workload c1r --levels-per-group 8"
for g in 2 8 16; do $B --bwd-levels-per-group $g > gpurun_out/ak_c1r_bg$g.json 2>> gpurun_out/ak.err; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/ak_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/ak.err
