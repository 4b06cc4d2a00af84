#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --skip-cpu > gpurun_out/r4d_bench_g2.json 2> gpurun_out/r4d_bench_g2.err; echo "bench_exit=$?"; tail -2 gpurun_out/r4d_bench_g2.err; cut -c1-600 gpurun_out/r4d_bench_g2.json
