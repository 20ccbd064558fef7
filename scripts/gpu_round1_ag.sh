#!/bin/bash
# experiment: forward table gathers as GPU-scope relaxed loads (DE_FLAG_COHERENT_GATHER 0x2000): does the 64 MiB c3 table group stay in L2?
set -x
mkdir -p gpurun_out
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B --workload c3 --flags 0x2000 > gpurun_out/ag_c3_coh.json 2>> gpurun_out/ag.err
$B --workload c3 --flags 0x2000 --levels-per-group 8 > gpurun_out/ag_c3_coh_g8.json 2>> gpurun_out/ag.err
$B --workload c2 --flags 0x2000 > gpurun_out/ag_c2_coh.json 2>> gpurun_out/ag.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/ag_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/ag.err
