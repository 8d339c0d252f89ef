#!/bin/bash
# Run every GPU test group in its own process (a faulting kernel must not poison the rest).
mkdir -p gpurun_out
for t in test_gemm add_rmsnorm rope_kv silu_mul "attn_decode" attn_prefill "sample"; do
  name=$(echo "$t" | tr ' ' '_')
  timeout 600 python -m pytest tests/test_kernels_gpu.py -q -k "$t" > "gpurun_out/k_${name}.log" 2>&1
  echo "kernels[$t] exit $?"; tail -3 "gpurun_out/k_${name}.log"
done
for t in "tiny" "eos" "llama1b"; do
  timeout 900 python -m pytest tests/test_engine_gpu.py -q -s -k "$t" > "gpurun_out/e_${t}.log" 2>&1
  echo "engine[$t] exit $?"; grep -E "^\[|passed|failed|Error|error" "gpurun_out/e_${t}.log" | tail -25
done
