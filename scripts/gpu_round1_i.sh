#!/bin/bash
# tcgen05 MLP head first run (own timeout: a wrong descriptor must not hang the box) + c3 uniform-vs-florida
set -x
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mlp_head or fused_head" > gpurun_out/pytest_mlp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mlp.log
tail -25 gpurun_out/pytest_mlp.log
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B --workload c3 --coords uniform > gpurun_out/i_c3_uniform.json 2>> gpurun_out/i.err
$B --workload c2 --coords florida > gpurun_out/i_c2_florida.json 2>> gpurun_out/i.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/i_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/i.err
