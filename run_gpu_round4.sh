#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q > gpurun_out/zz_alone.log 2>&1; echo "no_alloc alone exit $?"; tail -5 gpurun_out/zz_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/all_gpu.log 2>&1; echo "all gpu tests exit $?"; tail -8 gpurun_out/all_gpu.log
