#!/bin/bash
# ncu: launch list and full capture of the Mode R (c1r) kernels
set -x
mkdir -p gpurun_out
CMD="python bench.py --no-cpu-baseline --workload c1r --steps 2 --warmup 3"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_c1r_i.csv $CMD > gpurun_out/ncu_launches_c1r_i.log 2>&1
timeout 300 $CMD > gpurun_out/an_plain.json 2> gpurun_out/an.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"grid4d_(forward|backward)_kernel" -s 6 -c 4 -o gpurun_out/prof_c1r_i -f $CMD > gpurun_out/ncu_full_c1r_i.log 2>&1
ls -la gpurun_out/prof_c1r_i.ncu-rep; tail -3 gpurun_out/ncu_full_c1r_i.log
