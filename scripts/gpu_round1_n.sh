#!/bin/bash
# r01e evidence: GPU suite, bench lines (incl. packet roofline), coords / sort A-B, ncu launch list + full captures (post L2 hints)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/n_c2.json 2> gpurun_out/n.err
$B --workload c3 > gpurun_out/n_c3.json 2>> gpurun_out/n.err
$B --workload c5 > gpurun_out/n_c5.json 2>> gpurun_out/n.err
$B --workload c1r > gpurun_out/n_c1r.json 2>> gpurun_out/n.err
$B --workload c3 --coords uniform > gpurun_out/n_c3_uniform.json 2>> gpurun_out/n.err
$B --workload c3 --sort-points > gpurun_out/n_c3_sorted.json 2>> gpurun_out/n.err
$B --workload c2 --coords florida > gpurun_out/n_c2_florida.json 2>> gpurun_out/n.err
$B --workload c2 --coords florida --sort-points > gpurun_out/n_c2_florida_sorted.json 2>> gpurun_out/n.err
$B --workload c2 --sort-points > gpurun_out/n_c2_sorted.json 2>> gpurun_out/n.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/n_ref_c2.json 2>> gpurun_out/n.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_c2.csv $CMD > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:grid4d_ -s 6 -c 2 -o gpurun_out/prof_c2 -f $CMD > gpurun_out/ncu_full.log 2>&1
CMD3="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c3"
ncu --set full --clock-control none --import-source on -k regex:grid4d_ -s 6 -c 2 -o gpurun_out/prof_c3 -f $CMD3 > gpurun_out/ncu_full_c3.log 2>&1
CMD5="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c5 --log2-points 22"
ncu --set full --clock-control none --import-source on -k regex:mlp_head -s 6 -c 2 -o gpurun_out/prof_head -f $CMD5 > gpurun_out/ncu_full_head.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/n_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/n.err
ls -la gpurun_out | tail -30
