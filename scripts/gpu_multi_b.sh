#!/bin/bash
# multi-GPU weak scaling with the overlapped backward / all-reduce (default) and without (--no-overlap)
set -x
N=${1:-2}
mkdir -p gpurun_out
for w in c2 c3; do
for ov in "" "--no-overlap"; do
tag=$([ -z "$ov" ] && echo ov || echo plain)
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 --workload $w $ov > gpurun_out/n_${w}_n${N}_$tag.json 2> gpurun_out/n_${w}_n${N}_$tag.err
tail -2 gpurun_out/n_${w}_n${N}_$tag.err
done; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/n_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], round(d['value'],1), {k:round(v,2) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), d['clocks']['reasons'])
    except Exception as e: print(f, 'ERR', e)
PY
