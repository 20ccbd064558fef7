#!/bin/bash
# second GPU pass: gather peaks, per-level sweep, ncu launch list + full capture of fwd / bwd kernels
set -x
mkdir -p gpurun_out
timeout 600 python scripts/peaks_gather.py > gpurun_out/peaks_gather.log 2>&1; echo "rc=$?" >> gpurun_out/peaks_gather.log
timeout 600 python scripts/level_sweep.py --workload c2 > gpurun_out/level_sweep_c2.log 2>&1; echo "rc=$?" >> gpurun_out/level_sweep_c2.log
timeout 600 python scripts/level_sweep.py --workload c3 --log2-points 22 > gpurun_out/level_sweep_c3.log 2>&1; echo "rc=$?" >> gpurun_out/level_sweep_c3.log
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_c2.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:grid4d_ -s 6 -c 2 -o gpurun_out/prof_c2 -f $CMD > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
