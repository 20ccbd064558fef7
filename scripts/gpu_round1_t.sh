#!/bin/bash
# regrouped gradient rows (backward streams contiguous slices): suite + c3/c5/c2 A/B against in-place slices (--flags 0x800)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B --workload c3 > gpurun_out/t_c3.json 2>> gpurun_out/t.err
$B --workload c3 --flags 0x800 > gpurun_out/t_c3_noregroup.json 2>> gpurun_out/t.err
$B --workload c3 --bwd-levels-per-group 1 > gpurun_out/t_c3_bg1.json 2>> gpurun_out/t.err
$B --workload c3 --bwd-levels-per-group 4 > gpurun_out/t_c3_bg4.json 2>> gpurun_out/t.err
$B --workload c5 > gpurun_out/t_c5.json 2>> gpurun_out/t.err
$B --workload c2 > gpurun_out/t_c2.json 2>> gpurun_out/t.err
$B --workload c2 --bwd-levels-per-group 8 > gpurun_out/t_c2_bg8.json 2>> gpurun_out/t.err
CMD3="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c3"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"grid4d_backward|regroup" -s 6 -c 2 -o gpurun_out/prof_c3_regroup -f $CMD3 > gpurun_out/ncu_full_c3_regroup.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/t_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/t.err
