#!/bin/bash
# 8-GPU visit (charged 8x): multi-GPU parity subset, the bench under torchrun, halo-depth sweep.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "bands or shard or gpu_count or lifecycle_over" > gpurun_out/pytest_gpu8.log 2>&1; echo "pytest_exit=$?"
tail -2 gpurun_out/pytest_gpu8.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 8 --steps 2 --warmup 3 > gpurun_out/bench_g8.json 2> gpurun_out/bench_g8.err; echo "bench8_exit=$?"
ROWS=16384 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 \
    tools/conv_depth_sweep.py > gpurun_out/depth_rank_g8.log 2>&1; echo "sweep_rank_exit=$?"
ROWS=16384 python tools/conv_depth_sweep.py sp > gpurun_out/depth_sp_g8.log 2>&1; echo "sweep_sp_exit=$?"
grep -h "us/iter" gpurun_out/depth_rank_g8.log gpurun_out/depth_sp_g8.log | tail -40
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus 4 --steps 2 --warmup 3 > gpurun_out/bench_g4.json 2> gpurun_out/bench_g4.err; echo "bench4_exit=$?"
