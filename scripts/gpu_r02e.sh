#!/bin/bash
# round-2 GPU pass 2: tests, targeted bench lines, A/Bs, then the profiling pass of the new kernels
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -s 2>&1 | tail -60 > gpurun_out/r02e_pytest.txt
tail -4 gpurun_out/r02e_pytest.txt
for x in field projection_h768 st_embed; do
  timeout 200 python bench.py --extra-only $x --steps 3 > gpurun_out/r02e_$x.json 2> gpurun_out/r02e_$x.err || tail -3 gpurun_out/r02e_$x.err
  head -c 1500 gpurun_out/r02e_$x.json; echo
done
timeout 200 python bench.py --workload c1r --steps 10 --no-cpu-baseline > gpurun_out/r02e_c1r.json 2> gpurun_out/r02e_c1r.err
timeout 200 python bench.py --workload c1r --steps 10 --no-cpu-baseline --flags 0x4000 > gpurun_out/r02e_c1r_nosmall1d.json 2> gpurun_out/r02e_c1r_nosmall1d.err
timeout 300 python bench.py --workload c5h768 --steps 5 --no-cpu-baseline > gpurun_out/r02e_c5h768.json 2> gpurun_out/r02e_c5h768.err || tail -3 gpurun_out/r02e_c5h768.err
timeout 300 python scripts/ab_sorted.py > gpurun_out/r02e_ab_sorted.json 2> gpurun_out/r02e_ab_sorted.err || tail -3 gpurun_out/r02e_ab_sorted.err
python - <<'PY'
import json
for f in ("c1r", "c1r_nosmall1d", "c5h768"):
    try:
        d = json.load(open(f"gpurun_out/r02e_{f}.json"))
        print(f, round(d["value"], 1), d["kernels_ms"], "e2e", round(d["e2e"]["value"], 1))
    except Exception as e:
        print(f, "failed", e)
try:
    for c in json.load(open("gpurun_out/r02e_ab_sorted.json"))["cases"]:
        print(c)
except Exception as e:
    print("ab_sorted failed", e)
PY
bash scripts/gpu_r02d.sh proj field c1r > gpurun_out/r02e_prof.log 2>&1
tail -5 gpurun_out/r02e_prof.log
# last (a fault here must not cost the rest): the two-CTA multicast GEMM
DE_GEMM_MULTICAST=1 timeout 240 python -m pytest tests/test_projection.py -m gpu -q -s -x 2>&1 | tail -15 > gpurun_out/r02e_mc_pytest.txt
tail -4 gpurun_out/r02e_mc_pytest.txt
DE_GEMM_MULTICAST=1 timeout 200 python bench.py --extra-only projection_h768 --steps 3 > gpurun_out/r02e_mc_projection.json 2> gpurun_out/r02e_mc_projection.err || tail -3 gpurun_out/r02e_mc_projection.err
head -c 900 gpurun_out/r02e_mc_projection.json; echo
# two threads per row in the hidden kernel
DE_PROJ_HIDDEN_THREADS=256 timeout 240 python -m pytest tests/test_projection.py -m gpu -q -s -x -k "projection_forward or default_width" 2>&1 | tail -12 > gpurun_out/r02e_h256_pytest.txt
tail -3 gpurun_out/r02e_h256_pytest.txt
DE_PROJ_HIDDEN_THREADS=256 timeout 200 python bench.py --extra-only projection_h768 --steps 3 > gpurun_out/r02e_h256_projection.json 2> gpurun_out/r02e_h256_projection.err || tail -3 gpurun_out/r02e_h256_projection.err
head -c 300 gpurun_out/r02e_h256_projection.json; echo
