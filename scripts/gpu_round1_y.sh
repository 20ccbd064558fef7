#!/bin/bash
# host-side launch path trimmed (cached plans / packed levels, one working-table refresh): suite + bench lines
set -x
mkdir -p gpurun_out
echo skip-pytest

timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
B="timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
$B --workload c2 > gpurun_out/z_c2.json 2>> gpurun_out/z.err
$B --workload c3 > gpurun_out/z_c3.json 2>> gpurun_out/z.err
$B --workload c5 > gpurun_out/z_c5.json 2>> gpurun_out/z.err
$B --workload c1r > gpurun_out/z_c1r.json 2>> gpurun_out/z.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/z_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), 'launches', d['gpu_launches'])
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/z.err
