#!/bin/bash
# head backward ping-pong kernel: A/B against the serial kernel, sanitizer on failure
set -x
mkdir -p gpurun_out
for V in 0 2; do
  timeout 120 python scripts/head_bwd_check.py $V 128 1 129 5000 > gpurun_out/p_check_v$V.log 2>&1; echo "rc=$?" >> gpurun_out/p_check_v$V.log
  tail -5 gpurun_out/p_check_v$V.log
done
timeout 200 compute-sanitizer --tool memcheck --print-limit 20 python scripts/head_bwd_check.py 0 128 129 > gpurun_out/p_memcheck.log 2>&1; echo "rc=$?" >> gpurun_out/p_memcheck.log
grep -m 30 -A6 "Invalid\|Error\|ERROR SUMMARY" gpurun_out/p_memcheck.log | head -80
for V in 0 2; do
  timeout 120 python scripts/head_bwd_check.py $V 32 1 129 5000 75793 > gpurun_out/p_check32_v$V.log 2>&1; echo "rc=$?" >> gpurun_out/p_check32_v$V.log
  tail -5 gpurun_out/p_check32_v$V.log
  timeout 120 python scripts/head_bwd_check.py $V 128 75793 2097152 > gpurun_out/p_checkbig_v$V.log 2>&1; echo "rc=$?" >> gpurun_out/p_checkbig_v$V.log
  tail -3 gpurun_out/p_checkbig_v$V.log
done
