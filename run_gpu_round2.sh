#!/bin/bash
mkdir -p gpurun_out
for t in "tiny" "eos" "llama1b"; do
  timeout 900 python -m pytest tests/test_engine_gpu.py -q -s -k "$t" > "gpurun_out/e_${t}.log" 2>&1
  echo "engine[$t] exit $?"; grep -E "^\[|passed|failed|Error" "gpurun_out/e_${t}.log" | tail -25
done
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python scripts/profile_decode.py --steps 20 > gpurun_out/plain_decode.log 2>&1 && cat gpurun_out/plain_decode.log | tail -2 &&
timeout 300 python scripts/profile_decode.py --steps 20 --no-pdl 2>&1 | tail -1 &&
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_decode.csv python scripts/profile_decode.py --steps 2 > gpurun_out/ncu1.log 2>&1; echo "ncu1 exit $?"
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:attn_decode_kernel -c 2 -o gpurun_out/prof_attn_decode python scripts/profile_decode.py --steps 1 > gpurun_out/ncu2.log 2>&1; echo "ncu2 exit $?"
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_bf16 -c 5 -o gpurun_out/prof_gemm python scripts/profile_decode.py --steps 1 > gpurun_out/ncu3.log 2>&1; echo "ncu3 exit $?"
ls -la gpurun_out | tail -20
