#!/bin/bash
# round-2 GPU pass: established tests first, then the new tcgen05 / TMA kernels under their own short timeouts
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -s --ignore=tests/test_projection.py 2>&1 | tail -45 > gpurun_out/r02c_pytest.txt
tail -6 gpurun_out/r02c_pytest.txt
timeout 240 python -m pytest tests/test_projection.py -m gpu -q -s -x -k gemm_bf16_vs 2>&1 | tail -25 > gpurun_out/r02c_gemm.txt
tail -8 gpurun_out/r02c_gemm.txt
timeout 240 python -m pytest tests/test_projection.py -m gpu -q -s -x -k transposed 2>&1 | tail -25 > gpurun_out/r02c_gemm_t.txt
tail -8 gpurun_out/r02c_gemm_t.txt
timeout 240 python -m pytest tests/test_projection.py -m gpu -q -s -x -k projection_forward 2>&1 | tail -25 > gpurun_out/r02c_proj.txt
tail -8 gpurun_out/r02c_proj.txt
timeout 300 python -m pytest tests/test_projection.py -m gpu -q -s -x -k "backward or default_width" 2>&1 | tail -25 > gpurun_out/r02c_projb.txt
tail -8 gpurun_out/r02c_projb.txt
timeout 600 python bench.py --steps 5 > gpurun_out/r02c_bench.json 2> gpurun_out/r02c_bench.err
tail -2 gpurun_out/r02c_bench.err
python - <<'PY'
import json
try:
    d = json.load(open('gpurun_out/r02c_bench.json'))
    print('c2', d['value'], d['kernels_ms'])
    for k, v in d.get('extra', {}).items():
        if k in ('c3', 'c1r'):
            print(k, v['value'], v['kernels_ms'])
        else:
            print(k, {a: b for a, b in v.items() if a != 'what'} if isinstance(v, dict) else v)
except Exception as e:
    print('bench parse failed', e)
PY
