#!/bin/bash
# r01f evidence: suite, smoke, bench lines of every workload, reference arm, ncu launch lists (c2, c5) and head kernels
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/u_c2.json 2> gpurun_out/u.err
$B --workload c3 > gpurun_out/u_c3.json 2>> gpurun_out/u.err
$B --workload c5 > gpurun_out/u_c5.json 2>> gpurun_out/u.err
$B --workload c1r > gpurun_out/u_c1r.json 2>> gpurun_out/u.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/u_ref_c2.json 2>> gpurun_out/u.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload c5 --log2-points 22"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches_c5.csv $CMD > gpurun_out/ncu_launches_c5.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_head -s 6 -c 2 -o gpurun_out/prof_head5 -f $CMD > gpurun_out/ncu_full_head5.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/u_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3), {k:round(v['frac'],3) for k,v in d['roofline_packets'].items() if isinstance(v,dict)})
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/u.err; cat gpurun_out/smoke.log | tail -3
