#!/bin/bash
# c3 forward with 2 levels per group (32 MiB of tables per walk, 32 walks)
set -x
mkdir -p gpurun_out
B="timeout 400 python bench.py --no-cpu-baseline --workload c3 --steps 5 --warmup 3"
$B --levels-per-group 2 > gpurun_out/am_c3_g2.json 2>> gpurun_out/am.err
$B > gpurun_out/am_c3_g4.json 2>> gpurun_out/am.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/am_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/am.err
