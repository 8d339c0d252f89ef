for pm in "" 100 32; do
  for pl in 3 2; do
    if [ -z "$pm" ]; then CAPS=200,456,768,840,968 FT_PREFILL_PLANES=$pl timeout 100 python scripts/_mixed_time.py 2>&1 | grep MIXED
    else CAPS=200,456,768,840,968 FT_PREFILL_PLANES=$pl FT_GEMM_PAIR_MIN_TILES=$pm timeout 100 python scripts/_mixed_time.py 2>&1 | grep MIXED; fi
  done
done
