#!/bin/bash
# TMA-staged backward: parity, then A/B vs the register-prefetch pair kernel and ring-geometry sweeps
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/e_c2_staged.json 2>> gpurun_out/e.err
$B --flags 0x100 > gpurun_out/e_c2_nostage.json 2>> gpurun_out/e.err
for tp in 32 128 256; do DE_BWD_STAGE_POINTS=$tp $B > gpurun_out/e_c2_tp$tp.json 2>> gpurun_out/e.err; done
for st in 2 3 6; do DE_BWD_STAGES=$st $B > gpurun_out/e_c2_st$st.json 2>> gpurun_out/e.err; done
for cs in 4 128; do DE_BWD_CTA_STAGES=$cs $B > gpurun_out/e_c2_cs$cs.json 2>> gpurun_out/e.err; done
for bg in 4 8; do $B --bwd-levels-per-group $bg > gpurun_out/e_c2_bg$bg.json 2>> gpurun_out/e.err; done
$B --workload c3 > gpurun_out/e_c3_staged.json 2>> gpurun_out/e.err
$B --workload c3 --flags 0x100 > gpurun_out/e_c3_nostage.json 2>> gpurun_out/e.err
for bg in 1 4; do $B --workload c3 --bwd-levels-per-group $bg > gpurun_out/e_c3_bg$bg.json 2>> gpurun_out/e.err; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/e_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['clocks'])
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/e.err
