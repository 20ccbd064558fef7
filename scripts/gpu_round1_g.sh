#!/bin/bash
# defaults = plain lane-pair kernels with REDG: parity, level sweeps (contention on coarse levels), bench lines
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python scripts/level_sweep.py --workload c2 > gpurun_out/level_sweep_c2.log 2>&1
timeout 600 python scripts/level_sweep.py --workload c3 --log2-points 22 > gpurun_out/level_sweep_c3.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/g_c2.json 2> gpurun_out/g.err
tail -2 gpurun_out/level_sweep_c3.log; cat gpurun_out/g_c2.json
