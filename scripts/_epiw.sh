for W in 12 16; do
  sed -i "s/^#define FT_EPI_WARPS .*/#define FT_EPI_WARPS $W/" fasttalk_llm_microservice_b200/csrc/gemm.cu
  python -m fasttalk_llm_microservice_b200.build --force > /dev/null 2>&1 || { echo build failed; continue; }
  grep -A3 "gemm_pair_kernelILi128ELb0ELb1\|gemm_bf16_tcgen05_kernelILi64ELb1ELb0" fasttalk_llm_microservice_b200/_build/gemm.ptxas.log | grep "registers\|spill"
  echo "== W=$W cold prefill"; timeout 150 python scripts/timeline_prefill.py --sessions 4 --ctx 2048 2>&1 | grep -v Warn | grep "untraced\|gemm_pair\|attn_prefill_tc" 
  echo "== W=$W decode"; timeout 200 python bench.py --steps 20 --warmup 5 --no-ws --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['value'], d['e2e']['ttft_ms_p50'])"
  timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k gemm 2>&1 | tail -2
done
