#!/bin/bash
set -x
mkdir -p gpurun_out
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
timeout 300 python -m pytest tests -m gpu -x -q -k "mlp_head or fused_head" 2>&1 | tail -2
for S in 2 4; do
DE_HEAD_BWD_SPLIT=$S $B --workload c5 > gpurun_out/ac_c5_split$S.json 2>> gpurun_out/ac.err
done
CMD5="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c5 --log2-points 22"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_head -s 6 -c 2 -o gpurun_out/prof_head6 -f $CMD5 > gpurun_out/ncu_full_head6.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/ac_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/ac.err
