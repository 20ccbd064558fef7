#!/bin/bash
# small batches: step time of c2 / c1r at 2^12 .. 2^18 points (adaptive tile length), plus the suite
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/av_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/av_pytest.log
tail -2 gpurun_out/av_pytest.log
B="timeout 300 python bench.py --no-cpu-baseline --steps 50 --warmup 10"
for w in c2 c1r; do for lp in 12 14 16 18; do $B --workload $w --log2-points $lp > gpurun_out/av_${w}_p$lp.json 2>> gpurun_out/av.err; done; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/av_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['config']['points_per_gpu'], round(d['value'],1), 'ms/step', round(d['ms_per_step'],4), {k:round(v,4) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],4))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/av.err
