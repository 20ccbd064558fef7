#!/bin/bash
# forward loop trips per CTA, lane-pair kernels: c2 and c3
set -x
mkdir -p gpurun_out
B="timeout 400 python bench.py --no-cpu-baseline"
for it in 32 64 128 256; do DE_PAIR_TILE_ITERS_FWD=$it $B --workload c2 > gpurun_out/ap_c2_it$it.json 2>> gpurun_out/ap.err; done
for it in 8 32 64 128; do DE_PAIR_TILE_ITERS_FWD=$it $B --workload c3 --steps 5 --warmup 3 > gpurun_out/ap_c3_it$it.json 2>> gpurun_out/ap.err; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/ap_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/ap.err
