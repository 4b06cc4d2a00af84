#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/rank_mode_check.py > gpurun_out/r3m_rank8.out 2> gpurun_out/r3m_rank8.err; echo "rank8_exit=$?"
grep -v "^$" gpurun_out/r3m_rank8.out | tail -10; grep -i "error\|assert\|abandon\|Traceback" -A4 gpurun_out/r3m_rank8.err | head -60
TRI_GRID=crow timeout 300 python tools/nbody_tri_multi.py 1048576 8 > gpurun_out/r3m_multi_g8.txt 2>&1; echo "multi_exit=$?"; grep -v "^g=8 n=1048576 standard" gpurun_out/r3m_multi_g8.txt
