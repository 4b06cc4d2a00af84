#!/bin/bash
# 8 GPUs (charged 8x): the three pair-symmetric generations dealt over 8 GPUs, rank-mode parity under torchrun, bench at 8
mkdir -p gpurun_out
TRI_GRID=crow timeout 400 python tools/nbody_tri_multi.py 1048576 8 > gpurun_out/r3k_multi_g8.txt 2>&1; echo "multi_exit=$?"; cat gpurun_out/r3k_multi_g8.txt
timeout 600 python -m pytest tests/test_rank_mode_gpu.py -x -q -s -k "world_1 or 8" > gpurun_out/r3k_pytest_rank_g8.log 2>&1; echo "rank_exit=$?"; tail -12 gpurun_out/r3k_pytest_rank_g8.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 3 --warmup 3 --skip-cpu > gpurun_out/r3k_bench_g8.json 2> gpurun_out/r3k_bench_g8.err; echo "bench_exit=$?"; tail -3 gpurun_out/r3k_bench_g8.err; cut -c1-900 gpurun_out/r3k_bench_g8.json
