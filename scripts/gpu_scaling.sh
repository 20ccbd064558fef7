#!/bin/bash
# multi-GPU weak scaling: N ranks, points sharded, replicated tables, NCCL all-reduce of the flat gradient buffer
set -x
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
for w in c2 c3; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 --workload $w > gpurun_out/m_${w}_n$N.json 2> gpurun_out/m_${w}_n$N.err
tail -2 gpurun_out/m_${w}_n$N.err
done
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/m_c2_n1.json 2>> gpurun_out/m.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/m_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['clocks']['reasons'])
    except Exception as e: print(f, 'ERR', e)
PY
