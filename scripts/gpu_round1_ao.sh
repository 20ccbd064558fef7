#!/bin/bash
# loop trips per CTA (points per tile): c1r (point x level kernels) and c2 (lane-pair kernels)
set -x
mkdir -p gpurun_out
B="timeout 400 python bench.py --no-cpu-baseline"
for it in 2 8 16 32 64; do DE_TILE_ITERS=$it $B --workload c1r > gpurun_out/ao_c1r_it$it.json 2>> gpurun_out/ao.err; done
for it in 4 16 32; do DE_PAIR_TILE_ITERS=$it $B --workload c2 > gpurun_out/ao_c2_it$it.json 2>> gpurun_out/ao.err; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/ao_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/ao.err
