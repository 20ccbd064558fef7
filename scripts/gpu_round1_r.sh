#!/bin/bash
# dual-tile head forward + A&S GELU in the head kernels: suite, c5 A/B, ncu of the head kernels
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B --workload c5 > gpurun_out/s_c5.json 2>> gpurun_out/s.err
DE_HEAD_FWD_VARIANT=1 $B --workload c5 > gpurun_out/s_c5_fwd_single.json 2>> gpurun_out/s.err
CMD5="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c5 --log2-points 22"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_head -s 6 -c 2 -o gpurun_out/prof_head4 -f $CMD5 > gpurun_out/ncu_full_head4.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/s.err
