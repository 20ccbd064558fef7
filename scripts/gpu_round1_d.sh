#!/bin/bash
# pair-lane kernels (r01b): parity, smoke, bench c2 / c3 / c1r, gather peaks, ncu launch list + full capture
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "rc=$?" >> gpurun_out/bench_c2.err
timeout 600 python bench.py --steps 10 --warmup 3 --workload c1r --no-cpu-baseline > gpurun_out/bench_c1r.json 2> gpurun_out/bench_c1r.err; echo "rc=$?" >> gpurun_out/bench_c1r.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload c3 --no-cpu-baseline > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "rc=$?" >> gpurun_out/bench_c3.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "rc=$?" >> gpurun_out/bench_ref.err
timeout 600 python scripts/peaks_gather.py > gpurun_out/peaks_gather.log 2>&1; echo "rc=$?" >> gpurun_out/peaks_gather.log
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_c2.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:grid4d_ -s 6 -c 2 -o gpurun_out/prof_c2 -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench_c2.json gpurun_out/bench_c1r.json gpurun_out/bench_c3.json gpurun_out/bench_ref.json
ls -la gpurun_out
