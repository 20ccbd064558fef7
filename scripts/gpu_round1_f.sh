#!/bin/bash
# REDG (no-return) reductions + smem coarse forward + staged backward: parity, A/B matrix, ring sweeps
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
for fl in 0x0 0x100 0x200 0x300; do $B --flags $fl > gpurun_out/f_c2_fl$fl.json 2>> gpurun_out/f.err; done
for tp in 32 128; do DE_BWD_STAGE_POINTS=$tp $B > gpurun_out/f_c2_tp$tp.json 2>> gpurun_out/f.err; done
for bg in 4 8; do $B --bwd-levels-per-group $bg > gpurun_out/f_c2_bg$bg.json 2>> gpurun_out/f.err; $B --bwd-levels-per-group $bg --flags 0x100 > gpurun_out/f_c2_bg${bg}_ns.json 2>> gpurun_out/f.err; done
for fl in 0x0 0x100; do $B --workload c3 --flags $fl > gpurun_out/f_c3_fl$fl.json 2>> gpurun_out/f.err; done
for bg in 1 4; do $B --workload c3 --bwd-levels-per-group $bg --flags 0x100 > gpurun_out/f_c3_bg${bg}_ns.json 2>> gpurun_out/f.err; done
$B --workload c1r > gpurun_out/f_c1r.json 2>> gpurun_out/f.err
timeout 600 python scripts/peaks_gather.py > gpurun_out/peaks_gather.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/f_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['clocks']['sm_mhz'], d['clocks']['reasons'])
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/f.err
