#!/bin/bash
# round-2 third-generation kernel, first GPU visit: parity of nbody_crow_kernel, then timing sweeps
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_nbody_crow_gpu.py -x -q -k "not full_size" > gpurun_out/r3a_pytest_crow.log 2>&1; echo "pytest_exit=$?"; tail -6 gpurun_out/r3a_pytest_crow.log
timeout 400 python tools/nbody_crow_sweep.py 131072,262144 > gpurun_out/r3a_crow_sweep_small.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r3a_crow_sweep_small.txt
timeout 400 python tools/nbody_crow_sweep.py 1048576 3002:16384:8:32 3002:8192:4:16 3002:4096:2:8 3012:8192:4:16 3001:8192:4:16 3101:8192:4:16 3111:8192:4:16 > gpurun_out/r3a_crow_sweep_1m.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r3a_crow_sweep_1m.txt
