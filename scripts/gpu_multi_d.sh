#!/bin/bash
# N-GPU: grouped gradients through the overlapped level runs (c5), NCCL channel count / no-overlap A/B for c3 and c5
set -x
N=${1:-2}
mkdir -p gpurun_out
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3"
for w in c3 c5; do
$T --workload $w > gpurun_out/d_${w}_n${N}_ov.json 2> gpurun_out/d_${w}_n${N}_ov.err; tail -2 gpurun_out/d_${w}_n${N}_ov.err
$T --workload $w --no-overlap > gpurun_out/d_${w}_n${N}_plain.json 2> gpurun_out/d_${w}_n${N}_plain.err
NCCL_MAX_NCHANNELS=8 $T --workload $w > gpurun_out/d_${w}_n${N}_ov_ch8.json 2> gpurun_out/d_${w}_n${N}_ov_ch8.err
NCCL_MAX_NCHANNELS=4 $T --workload $w > gpurun_out/d_${w}_n${N}_ov_ch4.json 2> gpurun_out/d_${w}_n${N}_ov_ch4.err
$T --workload $w --sync-chunks 8 > gpurun_out/d_${w}_n${N}_ov_k8.json 2> gpurun_out/d_${w}_n${N}_ov_k8.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/d_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['config'].get('head_grad_group_width'))
    except Exception as e: print(f, 'ERR', e)
PY
