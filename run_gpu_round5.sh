#!/bin/bash
./run_gpu_tests.sh
timeout 300 python scripts/trace_gemm.py 2>&1 | tail -12
timeout 300 python scripts/bench_gemm.py 2>&1 | head -12
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-700 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python scripts/profile_decode.py --steps 20 > gpurun_out/plain_decode.log 2>&1 && tail -1 gpurun_out/plain_decode.log &&
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_decode.csv python scripts/profile_decode.py --steps 2 > gpurun_out/ncu1.log 2>&1; echo "ncu1 exit $?"
