#!/bin/bash
set -x
mkdir -p gpurun_out
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
for S in 1 2 3; do
DE_HEAD_BWD_SPLIT=$S timeout 200 python -m pytest tests -m gpu -x -q -k "mlp_head or fused_head" 2>&1 | tail -2
DE_HEAD_BWD_SPLIT=$S $B --workload c5 > gpurun_out/ab_c5_split$S.json 2>> gpurun_out/ab.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/ab_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/ab.err
