#!/bin/bash
# backward loop trips per CTA, lane-pair kernels, c3 (2 levels per group)
set -x
mkdir -p gpurun_out
B="timeout 400 python bench.py --no-cpu-baseline --workload c3 --steps 5 --warmup 3"
for it in 2 4 8 16 32; do DE_PAIR_TILE_ITERS_BWD=$it $B > gpurun_out/ar_c3_bit$it.json 2>> gpurun_out/ar.err; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/ar_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/ar.err
