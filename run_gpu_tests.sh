#!/bin/bash
# Run every GPU test group in its own process (a faulting kernel must not poison the rest).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
for t in test_gemm_tcgen05 test_gemm_bf16_out test_gemm_linearity add_rmsnorm rope_kv silu_mul "attn_decode and not cfg2" attn_decode_cfg2 attn_prefill "sample"; do
  name=$(echo "$t" | tr ' ' '_')
  timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k "$t" > "gpurun_out/k_${name}.log" 2>&1
  echo "kernels[$t] exit $?"
  tail -3 "gpurun_out/k_${name}.log"
done
for t in "tiny" "eos" "llama1b"; do
  timeout 900 python -m pytest tests/test_engine_gpu.py -x -q -s -k "$t" > "gpurun_out/e_${t}.log" 2>&1
  echo "engine[$t] exit $?"
  tail -5 "gpurun_out/e_${t}.log"
done
