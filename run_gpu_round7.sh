#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_engine_gpu.py -q -s -k larger > gpurun_out/e_larger.log 2>&1; echo "larger exit $?"; grep -E "^\[|passed|failed|Error" gpurun_out/e_larger.log | tail
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench2 exit $?"; cut -c1-600 gpurun_out/bench2.json; tail -3 gpurun_out/bench2.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 | cut -c1-500
