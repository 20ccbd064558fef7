#!/bin/bash
# r01i evidence: suite, smoke, bench lines of every workload with the new tile defaults, c2 launch list
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/aq_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/aq_pytest.log
tail -3 gpurun_out/aq_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/aq_smoke.log 2>&1; tail -1 gpurun_out/aq_smoke.log
timeout 400 python bench.py > gpurun_out/r01i_bench_c2.json 2>> gpurun_out/aq.err
timeout 400 python bench.py --workload c1r --no-cpu-baseline > gpurun_out/r01i_bench_c1r.json 2>> gpurun_out/aq.err
timeout 400 python bench.py --workload c3 --no-cpu-baseline > gpurun_out/r01i_bench_c3.json 2>> gpurun_out/aq.err
timeout 400 python bench.py --workload c5 --no-cpu-baseline > gpurun_out/r01i_bench_c5.json 2>> gpurun_out/aq.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r01i_launches_c2.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/aq_ncu.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r01i_bench_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), {k:round(v,3) for k,v in d['kernels_ms'].items()}, 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3), round(d['roofline_packets']['forward']['frac'],3), round(d['roofline_packets']['backward']['frac'],3))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/aq.err
