#!/bin/bash
# gradient replicas for contended coarse levels: parity + A/B on c2 / c3 / c1r
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
for w in c2 c3 c1r; do
$B --workload $w > gpurun_out/h_${w}_rep.json 2>> gpurun_out/h.err
$B --workload $w --no-replicas > gpurun_out/h_${w}_norep.json 2>> gpurun_out/h.err
done
timeout 600 python scripts/level_sweep.py --workload c3 --log2-points 24 > gpurun_out/level_sweep_c3.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/h_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['clocks']['sm_mhz'], d['clocks']['reasons'])
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/h.err
