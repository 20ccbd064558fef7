#!/bin/bash
# r01c evidence: full GPU suite, bench lines, ncu launch list + full captures for c2 and c3 (default kernels)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/j_c2.json 2> gpurun_out/j.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload c3 --no-cpu-baseline > gpurun_out/j_c3.json 2>> gpurun_out/j.err
timeout 600 python bench.py --steps 10 --warmup 3 --workload c1r --no-cpu-baseline > gpurun_out/j_c1r.json 2>> gpurun_out/j.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_c2.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:grid4d_ -s 6 -c 2 -o gpurun_out/prof_c2 -f $CMD > gpurun_out/ncu_full.log 2>&1
CMD3="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload c3"
$CMD3 > gpurun_out/plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:grid4d_ -s 6 -c 2 -o gpurun_out/prof_c3 -f $CMD3 > gpurun_out/ncu_full_c3.log 2>&1
ls -la gpurun_out | tail -20
