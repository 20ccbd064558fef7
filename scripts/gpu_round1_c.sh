#!/bin/bash
# pair-lane kernels: parity + bench sweeps over levels-per-group and L2 fetch granularity
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for g in 4 8 16; do for bg in 4 8 16; do
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --levels-per-group $g --bwd-levels-per-group $bg > gpurun_out/bench_c2_g${g}_bg${bg}.json 2>> gpurun_out/bench_sweep.err
done; done
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --levels-per-group 4 --bwd-levels-per-group 4 --l2-fetch 32 > gpurun_out/bench_c2_g4_bg4_f32.json 2>> gpurun_out/bench_sweep.err
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --l2-fetch 32 > gpurun_out/bench_c2_auto_f32.json 2>> gpurun_out/bench_sweep.err
for bg in 1 2 4; do
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload c3 --levels-per-group 4 --bwd-levels-per-group $bg > gpurun_out/bench_c3_g4_bg${bg}.json 2>> gpurun_out/bench_sweep.err
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload c3 --levels-per-group 4 --bwd-levels-per-group $bg --l2-fetch 32 > gpurun_out/bench_c3_g4_bg${bg}_f32.json 2>> gpurun_out/bench_sweep.err
done
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload c3 --levels-per-group 8 --bwd-levels-per-group 2 --l2-fetch 32 > gpurun_out/bench_c3_g8_bg2_f32.json 2>> gpurun_out/bench_sweep.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_c[23]_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['value'],1), d['kernels_ms'], 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
