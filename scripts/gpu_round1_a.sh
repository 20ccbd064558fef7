#!/bin/bash
# first GPU pass: parity tests, smoke, bench on c2 / c1r / c3
set -x
mkdir -p gpurun_out
nvidia-smi > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "rc=$?" >> gpurun_out/bench_c2.err
timeout 600 python bench.py --steps 10 --warmup 3 --workload c1r --no-cpu-baseline > gpurun_out/bench_c1r.json 2> gpurun_out/bench_c1r.err; echo "rc=$?" >> gpurun_out/bench_c1r.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload c3 --no-cpu-baseline > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "rc=$?" >> gpurun_out/bench_c3.err
tail -3 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench_c2.json gpurun_out/bench_c2.err gpurun_out/bench_c1r.json gpurun_out/bench_c3.json gpurun_out/bench_c3.err
