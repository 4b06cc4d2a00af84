#!/bin/bash
# 2 GPUs: rank-mode (torchrun) parity incl. the pair-symmetric wavefront over CUDA IPC, the single-process
# multi-GPU tests, and bench.py under torchrun
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_rank_mode_gpu.py -x -q -s > gpurun_out/r2k_pytest_rank.log 2>&1; echo "rank_exit=$?"; tail -8 gpurun_out/r2k_pytest_rank.log
timeout 900 python -m pytest tests/test_nbody_tri_gpu.py tests/test_nbody_sym_gpu.py -x -q > gpurun_out/r2k_pytest_tri.log 2>&1; echo "tri_exit=$?"; tail -4 gpurun_out/r2k_pytest_tri.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --skip-cpu > gpurun_out/r2k_bench_g2.json 2> gpurun_out/r2k_bench_g2.err; echo "bench_exit=$?"; tail -3 gpurun_out/r2k_bench_g2.err; cut -c1-1500 gpurun_out/r2k_bench_g2.json
