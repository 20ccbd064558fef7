#!/bin/bash
# round-2 GPU pass: tests + smoke, then the bench lines of the reworked kernels and a launch list of the projection
mkdir -p gpurun_out
T=${1:-r02x}
timeout 1200 python -m pytest tests -m gpu -q -s 2>&1 | tail -30 > gpurun_out/${T}_pytest.txt
tail -3 gpurun_out/${T}_pytest.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
for x in field projection_h768; do
  timeout 200 python bench.py --extra-only $x --steps 3 > gpurun_out/${T}_$x.json 2> gpurun_out/${T}_$x.err || tail -3 gpurun_out/${T}_$x.err
done
DE_GEMM_PAIR=0 timeout 200 python bench.py --extra-only projection_h768 --steps 3 > gpurun_out/${T}_projection_h768_single_cta.json 2>/dev/null
timeout 300 python bench.py --workload c5h768 --steps 5 --no-cpu-baseline > gpurun_out/${T}_c5h768.json 2> gpurun_out/${T}_c5h768.err || tail -3 gpurun_out/${T}_c5h768.err
python - $T <<'PY'
import json, sys
T = sys.argv[1]
for f in ("projection_h768", "projection_h768_single_cta"):
    try:
        d = json.load(open(f"gpurun_out/{T}_{f}.json"))
        print(f, "fwd", round(d["forward_ms"], 3), "fwd+bwd", round(d["fwd_bwd_ms"], 3), "gemm2 fp32", round(d["gemm2"]["ms_fp32_out"], 3),
              "bf16", round(d["gemm2"]["ms_bf16_out"], 3), "frac sustained", round(d["gemm2"]["frac_of_bf16_sustained"], 3))
    except Exception as e:
        print(f, "failed", e)
try:
    d = json.load(open(f"gpurun_out/{T}_field.json"))
    print("field", round(d["field_forward_ms"], 3), "general", round(d["general_forward_ms"], 3), d["bit_identical"])
    d = json.load(open(f"gpurun_out/{T}_c5h768.json"))
    print("c5h768", round(d["value"], 1), d["kernels_ms"], "e2e", round(d["e2e"]["value"], 1))
except Exception as e:
    print("failed", e)
PY
CMD="python bench.py --extra-only projection_h768 --steps 2"
$CMD > gpurun_out/${T}_plain_proj.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches_proj.csv $CMD > gpurun_out/${T}_ncu1_proj.log 2>&1
python scripts/summarize_profile.py launches gpurun_out/${T}_launches_proj.csv 2>/dev/null | grep "de::" | cut -c1-140
