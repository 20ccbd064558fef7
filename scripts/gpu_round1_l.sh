#!/bin/bash
# tcgen05 head backward first run (own timeout) + c5 bench
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mlp_head or fused_head" > gpurun_out/pytest_mlp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mlp.log
tail -30 gpurun_out/pytest_mlp.log
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B --workload c5 --log2-points 22 > gpurun_out/l_c5_22.json 2>> gpurun_out/l.err
$B --workload c5 > gpurun_out/l_c5.json 2>> gpurun_out/l.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/l_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/l.err
