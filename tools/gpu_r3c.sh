#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_nbody_crow_gpu.py tests/test_nbody_tri_gpu.py -x -q -k "every_variant or ragged" > gpurun_out/r3c_pytest.log 2>&1; echo "pytest_exit=$?"; tail -3 gpurun_out/r3c_pytest.log
timeout 400 python tools/nbody_crow_sweep.py 65536,131072 3002:16384:8:32 3002:4096:4:16 3002:2048:4:16 3012:16384:8:32 > gpurun_out/r3c_crow_sweep_small.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r3c_crow_sweep_small.txt
timeout 400 python tools/nbody_crow_sweep.py 1048576 3002:16384:8:32 3002:8192:4:16 3002:4096:2:8 3002:32768:8:32 3012:8192:4:16 > gpurun_out/r3c_crow_sweep_1m.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r3c_crow_sweep_1m.txt
