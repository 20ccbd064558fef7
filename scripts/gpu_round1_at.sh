#!/bin/bash
# replica cap per level group: suite, c4 at N=1, c2 / c3 unchanged?
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/at_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/at_pytest.log
tail -3 gpurun_out/at_pytest.log
B="timeout 600 python bench.py --no-cpu-baseline"
$B --workload c4 --steps 5 --warmup 3 > gpurun_out/r01i_bench_c4_n1.json 2> gpurun_out/at.err
$B --workload c2 > gpurun_out/at_c2.json 2>> gpurun_out/at.err
$B --workload c3 --steps 5 --warmup 3 > gpurun_out/at_c3.json 2>> gpurun_out/at.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r01i_bench_c4*.json'))+sorted(glob.glob('gpurun_out/at_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], d['scaling'], d['config']['points_per_gpu'], round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/at.err
