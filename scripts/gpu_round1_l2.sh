#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "mlp_head or fused_head" > gpurun_out/pytest_mlp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mlp.log
grep -E "assert|Error|passed|failed" gpurun_out/pytest_mlp.log | tail -30
