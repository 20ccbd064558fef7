#!/bin/bash
# Mode R (c1r): levels per group sweep (lanes of a point = levels of its group)
set -x
mkdir -p gpurun_out
B="timeout 400 python bench.py --no-cpu-baseline --workload c1r"
for g in 2 4 8 16; do $B --levels-per-group $g > gpurun_out/aj_c1r_g$g.json 2>> gpurun_out/aj.err; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/aj_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/aj.err
