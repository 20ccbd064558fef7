#!/bin/bash
# end-of-session verification of the committed tree (ABI v3): GPU tests, smoke, default bench, reference arm
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/ah_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ah_pytest.log
tail -3 gpurun_out/ah_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/ah_smoke.log 2>&1; tail -2 gpurun_out/ah_smoke.log
timeout 400 python bench.py > gpurun_out/ah_bench_c2.json 2> gpurun_out/ah.err; tail -c 600 gpurun_out/ah_bench_c2.json
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/ah_bench_ref.json 2>> gpurun_out/ah.err; tail -c 400 gpurun_out/ah_bench_ref.json
tail -3 gpurun_out/ah.err
