#!/bin/bash
# r01f multi-GPU check after the regroup pass / head kernels: c2, c3, c5 weak scaling at N GPUs (overlapped all-reduce)
set -x
N=${1:-2}
mkdir -p gpurun_out
for w in c2 c3 c5; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 --workload $w > gpurun_out/v_${w}_n${N}.json 2> gpurun_out/v_${w}_n${N}.err
tail -2 gpurun_out/v_${w}_n${N}.err
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 2 --warmup 1 --impl reference > gpurun_out/v_ref_n${N}.json 2> gpurun_out/v_ref_n${N}.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/v_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['clocks']['reasons'])
    except Exception as e: print(f, 'ERR', e)
PY
cat gpurun_out/v_ref_n${N}.json | cut -c1-200
