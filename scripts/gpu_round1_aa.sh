#!/bin/bash
# head backward register split A/B (MMA warpgroup 56 / consumers 224 vs 88 / 208), c5
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q -k "mlp_head or fused_head" 2>&1 | tail -3
DE_HEAD_BWD_SPLIT=1 timeout 300 python -m pytest tests -m gpu -x -q -k "mlp_head or fused_head" 2>&1 | tail -3
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B --workload c5 > gpurun_out/aa_c5_split0.json 2>> gpurun_out/aa.err
DE_HEAD_BWD_SPLIT=1 $B --workload c5 > gpurun_out/aa_c5_split1.json 2>> gpurun_out/aa.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/aa_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/aa.err
