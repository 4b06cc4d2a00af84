#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/rank_mode_check.py > gpurun_out/r3l_rank2.out 2> gpurun_out/r3l_rank2.err; echo "rank2_exit=$?"
grep -v "^$" gpurun_out/r3l_rank2.out | tail -5; grep -i "error\|assert\|abandon\|Traceback" -A3 gpurun_out/r3l_rank2.err | head -40
