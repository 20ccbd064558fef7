#!/bin/bash
# ping-pong head backward (mlp_head_bwd2.cu) parity + A/B; c3 L2 residency experiments (persisting set-aside, fetch granularity, group sizes)
set -x
mkdir -p gpurun_out
for V in 0 2 1; do
  DE_HEAD_BWD_VARIANT=$V timeout 300 python -m pytest tests -m gpu -x -q -k "mlp_head or fused_head" > gpurun_out/o_head_v$V.log 2>&1; echo "pytest rc=$?" >> gpurun_out/o_head_v$V.log
  tail -4 gpurun_out/o_head_v$V.log
done
B="timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
for V in 0 2 1; do
  DE_HEAD_BWD_VARIANT=$V $B --workload c5 > gpurun_out/o_c5_v$V.json 2>> gpurun_out/o.err
done
$B --workload c3 --persist-l2-mb -1 > gpurun_out/o_c3_persist.json 2>> gpurun_out/o.err
$B --workload c3 --l2-fetch 32 > gpurun_out/o_c3_f32.json 2>> gpurun_out/o.err
$B --workload c3 --persist-l2-mb -1 --l2-fetch 32 > gpurun_out/o_c3_persist_f32.json 2>> gpurun_out/o.err
$B --workload c3 --persist-l2-mb -1 --bwd-levels-per-group 4 --levels-per-group 8 > gpurun_out/o_c3_persist_g8_bg4.json 2>> gpurun_out/o.err
$B --workload c3 --persist-l2-mb -1 --bwd-levels-per-group 1 --levels-per-group 2 > gpurun_out/o_c3_persist_g2_bg1.json 2>> gpurun_out/o.err
$B --workload c3 --persist-l2-mb 32 > gpurun_out/o_c3_persist32.json 2>> gpurun_out/o.err
$B --workload c3 --persist-l2-mb 64 > gpurun_out/o_c3_persist64.json 2>> gpurun_out/o.err
$B --workload c2 --persist-l2-mb -1 > gpurun_out/o_c2_persist.json 2>> gpurun_out/o.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/o_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['config'].get('l2_persisting_bytes'), d['config'].get('l2_persisting_max_bytes'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/o.err
