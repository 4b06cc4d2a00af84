#!/bin/bash
mkdir -p gpurun_out
HGB_NBODY_TRI_FLAGS=3 timeout 300 python -m pytest tests/test_nbody_crow_gpu.py -x -q -k "every_variant or ragged or dealt" > gpurun_out/r3f_pytest.log 2>&1; echo "pytest_exit=$?"; tail -3 gpurun_out/r3f_pytest.log
timeout 400 python tools/nbody_crow_sweep.py 131072 3002:16384:8:32:0 3002:16384:8:32:1 3002:16384:8:32:2 3002:16384:8:32:3 3012:16384:8:32:3 > gpurun_out/r3f_crow_sweep_small.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r3f_crow_sweep_small.txt
timeout 400 python tools/nbody_crow_sweep.py 1048576 3002:4096:2:8:0 3002:4096:2:8:1 3002:4096:2:8:2 3002:4096:2:8:3 3002:16384:8:32:3 3012:4096:2:8:3 > gpurun_out/r3f_crow_sweep_1m.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r3f_crow_sweep_1m.txt
