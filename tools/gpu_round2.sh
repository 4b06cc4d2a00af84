#!/bin/bash
# 2-GPU visit: GPU tests (real NCCL + peer paths), the bench under torchrun, the halo-depth sweep.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest_exit=$?"
tail -3 gpurun_out/pytest_gpu2.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/bench_g2.json 2> gpurun_out/bench_g2.err; echo "bench2_exit=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
    tools/conv_depth_sweep.py > gpurun_out/depth_rank_g2.log 2>&1; echo "sweep_rank_exit=$?"
python tools/conv_depth_sweep.py sp > gpurun_out/depth_sp_g2.log 2>&1; echo "sweep_sp_exit=$?"
grep -h "us/iter" gpurun_out/depth_rank_g2.log gpurun_out/depth_sp_g2.log | tail -60
