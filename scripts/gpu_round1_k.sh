#!/bin/bash
# L2 eviction-policy hints (evict_last tables / evict_first streams): parity + A/B; c5 first run
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
for w in c2 c3 c1r; do
$B --workload $w > gpurun_out/k_${w}_hints.json 2>> gpurun_out/k.err
$B --workload $w --flags 0x400 > gpurun_out/k_${w}_nohints.json 2>> gpurun_out/k.err
done
for bg in 1 4; do $B --workload c3 --bwd-levels-per-group $bg > gpurun_out/k_c3_bg$bg.json 2>> gpurun_out/k.err; done
for g in 2 8; do $B --workload c3 --levels-per-group $g > gpurun_out/k_c3_g$g.json 2>> gpurun_out/k.err; done
$B --workload c5 --log2-points 22 > gpurun_out/k_c5_22.json 2>> gpurun_out/k.err
$B --workload c5 > gpurun_out/k_c5.json 2>> gpurun_out/k.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/k_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['clocks']['reasons'])
    except Exception as e: print(f, 'ERR', e)
PY
tail -8 gpurun_out/k.err
