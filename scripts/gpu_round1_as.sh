#!/bin/bash
# c4 (2^26 points in total) at N = 1 and N = 2
set -x
mkdir -p gpurun_out
N=${1:-1}
if [ "$N" = 1 ]; then
  timeout 600 python bench.py --workload c4 --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/r01i_bench_c4_n1.json 2> gpurun_out/as.err
else
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --workload c4 --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/r01i_bench_c4_n$N.json 2> gpurun_out/as.err
fi
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r01i_bench_c4*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], d['scaling'], d['config']['points_per_gpu'], round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/as.err
