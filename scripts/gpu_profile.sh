#!/bin/bash
# One gpurun call's worth of evidence for profiles/: a plain run first (its exit code gates the rest),
# then the ncu launch list of two decode steps, full captures of the decode-attention kernel, the
# decode GEMMs and the persistent prefill GEMMs, and the CUPTI timelines (real PDL overlap).
#   gpurun --timeout 1800 -- scripts/gpu_profile.sh      -> gpurun_out/
mkdir -p gpurun_out
timeout 300 python scripts/profile_decode.py --steps 20 > gpurun_out/plain_decode.log 2>&1 && tail -1 gpurun_out/plain_decode.log || exit 1
timeout 300 python scripts/timeline_prefill.py --sessions 4 --ctx 2048 --no-trace > gpurun_out/plain_prefill.log 2>&1 && tail -1 gpurun_out/plain_prefill.log || exit 1
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_decode.csv python scripts/profile_decode.py --steps 2 > gpurun_out/ncu1.log 2>&1; echo "ncu1 exit $?"
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:attn_decode_kernel -c 2 -o gpurun_out/prof_attn_decode -f python scripts/profile_decode.py --steps 1 > gpurun_out/ncu2.log 2>&1; echo "ncu2 exit $?"
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_ -c 6 -o gpurun_out/prof_gemm -f python scripts/profile_decode.py --steps 1 > gpurun_out/ncu3.log 2>&1; echo "ncu3 exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_ -s 130 -c 5 -o gpurun_out/prof_gemm_prefill -f python scripts/timeline_prefill.py --sessions 4 --ctx 2048 --no-trace > gpurun_out/ncu4.log 2>&1; echo "ncu4 exit $?"
timeout 300 python scripts/timeline_decode.py > gpurun_out/timeline.txt 2>&1; tail -12 gpurun_out/timeline.txt
timeout 300 python scripts/timeline_prefill.py --sessions 4 --ctx 2048 > gpurun_out/timeline_prefill.txt 2>&1; grep -v Warn gpurun_out/timeline_prefill.txt | tail -22
