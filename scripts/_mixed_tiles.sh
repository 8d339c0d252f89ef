timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k "fp16_single_plane or cta_pair" 2>&1 | tail -5
echo "== mixed tiles"; timeout 150 python scripts/timeline_prefill.py --sessions 4 --ctx 2048 2>&1 | grep -v Warn | grep "untraced\|gemm_pair\|attn_prefill_tc" | head -14
echo "== no mixed"; FT_GEMM_NO_MIXED_TILES=1 timeout 150 python scripts/timeline_prefill.py --sessions 4 --ctx 2048 2>&1 | grep -v Warn | grep "untraced\|gemm_pair<" | head -4
timeout 400 python -m pytest tests/test_engine_gpu.py -x -q 2>&1 | tail -3
