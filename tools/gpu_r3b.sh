#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_nbody_crow_gpu.py -x -q -k "every_variant or ragged" > gpurun_out/r3b_pytest_crow.log 2>&1; echo "pytest_exit=$?"; tail -3 gpurun_out/r3b_pytest_crow.log
timeout 400 python tools/nbody_crow_sweep.py 131072,262144 3002:16384:8:32 3002:8192:8:32 3002:4096:4:16 3002:4096:2:8 3012:8192:8:32 3001:8192:4:16 3011:8192:4:16 3101:8192:4:16 3111:8192:4:16 > gpurun_out/r3b_crow_sweep_small.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r3b_crow_sweep_small.txt
timeout 400 python tools/nbody_crow_sweep.py 1048576 3002:16384:8:32 3002:8192:4:16 3012:8192:4:16 3001:8192:4:16 3101:8192:4:16 > gpurun_out/r3b_crow_sweep_1m.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r3b_crow_sweep_1m.txt
