#!/bin/bash
# round-2 GPU pass 3: tests, then the reworked projection / field kernels (bench lines + launch list + full ncu set)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -s 2>&1 | tail -40 > gpurun_out/r02g_pytest.txt
tail -4 gpurun_out/r02g_pytest.txt
for x in field projection_h768; do
  timeout 200 python bench.py --extra-only $x --steps 3 > gpurun_out/r02g_$x.json 2> gpurun_out/r02g_$x.err || tail -3 gpurun_out/r02g_$x.err
  head -c 420 gpurun_out/r02g_$x.json; echo
done
DE_PROJ_HIDDEN_THREADS=128 timeout 200 python bench.py --extra-only projection_h768 --steps 3 > gpurun_out/r02g_projection_h768_t128.json 2>/dev/null
head -c 200 gpurun_out/r02g_projection_h768_t128.json; echo
timeout 300 python bench.py --workload c5h768 --steps 5 --no-cpu-baseline > gpurun_out/r02g_c5h768.json 2> gpurun_out/r02g_c5h768.err || tail -3 gpurun_out/r02g_c5h768.err
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/r02g_c5h768.json"))
    print("c5h768", round(d["value"], 1), d["kernels_ms"], "e2e", round(d["e2e"]["value"], 1))
except Exception as e:
    print("c5h768 failed", e)
PY
sed -i 's/r02f_/r02g_/g' scripts/gpu_r02d.sh
bash scripts/gpu_r02d.sh proj field > gpurun_out/r02g_prof.log 2>&1
tail -3 gpurun_out/r02g_prof.log
DE_GEMM_MULTICAST=1 timeout 200 python bench.py --extra-only projection_h768 --steps 3 > gpurun_out/r02g_projection_h768_mc.json 2>/dev/null
DE_GEMM_PREFETCH=0 timeout 200 python bench.py --extra-only projection_h768 --steps 3 > gpurun_out/r02g_projection_h768_nopf.json 2>/dev/null
python - <<'PY'
import json
for f in ("projection_h768", "projection_h768_nopf", "projection_h768_t128", "projection_h768_mc"):
    try:
        d = json.load(open(f"gpurun_out/r02g_{f}.json"))
        print(f, "fwd", round(d["forward_ms"], 3), "fwd+bwd", round(d["fwd_bwd_ms"], 3), "gemm2 fp32", round(d["gemm2"]["ms_fp32_out"], 3), "bf16", round(d["gemm2"]["ms_bf16_out"], 3))
    except Exception as e:
        print(f, "failed", e)
PY
