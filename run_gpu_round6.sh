#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/all_gpu.log 2>&1; echo "all gpu tests exit $?"; tail -4 gpurun_out/all_gpu.log
timeout 300 python scripts/timeline_decode.py 2>&1 | tail -34
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-330 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
