#!/bin/bash
# r01h evidence: suite, smoke, bench lines of every workload (default line with cpu baseline), reference arm,
# ncu launch lists (c2 default command, c5) and full captures of the head kernels
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/h_c2.json 2> gpurun_out/h.err
B="timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
$B --workload c3 > gpurun_out/h_c3.json 2>> gpurun_out/h.err
$B --workload c5 > gpurun_out/h_c5.json 2>> gpurun_out/h.err
$B --workload c1r > gpurun_out/h_c1r.json 2>> gpurun_out/h.err
$B --workload c3 --coords uniform > gpurun_out/h_c3_uniform.json 2>> gpurun_out/h.err
$B --workload c2 --coords florida > gpurun_out/h_c2_florida.json 2>> gpurun_out/h.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/h_ref_c2.json 2>> gpurun_out/h.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_c2_h.csv $CMD > gpurun_out/ncu_launches_h.log 2>&1
CMD5="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload c5 --log2-points 22"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches_c5_h.csv $CMD5 > gpurun_out/ncu_launches_c5_h.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_head -s 6 -c 2 -o gpurun_out/prof_head_h -f $CMD5 > gpurun_out/ncu_full_head_h.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/h_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3), {k:round(v['frac'],3) for k,v in d['roofline_packets'].items() if isinstance(v,dict)}, 'launches', d['gpu_launches'])
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/h.err
