#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/rank_mode_check.py > gpurun_out/r4b_rank2.out 2> gpurun_out/r4b_rank2.err; echo "rank2_exit=$?"; grep "RANK_MODE_OK\|cases ran" gpurun_out/r4b_rank2.out; tail -2 gpurun_out/r4b_rank2.err
TRI_GRID=seg timeout 400 python tools/nbody_tri_multi.py 1048576 2 > gpurun_out/r4b_multi_g2.txt 2>&1; echo "multi_exit=$?"; cat gpurun_out/r4b_multi_g2.txt
