#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/profile_decode.py --steps 20 > gpurun_out/plain_decode.log 2>&1 && tail -1 gpurun_out/plain_decode.log &&
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_decode.csv python scripts/profile_decode.py --steps 2 > gpurun_out/ncu1.log 2>&1; echo "ncu1 exit $?"
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:attn_decode_kernel -c 2 -o gpurun_out/prof_attn_decode python scripts/profile_decode.py --steps 1 > gpurun_out/ncu2.log 2>&1; echo "ncu2 exit $?"
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_bf16 -c 5 -o gpurun_out/prof_gemm python scripts/profile_decode.py --steps 1 > gpurun_out/ncu3.log 2>&1; echo "ncu3 exit $?"
timeout 300 python scripts/timeline_decode.py > gpurun_out/timeline.txt 2>&1; tail -12 gpurun_out/timeline.txt
