#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_nbody_crow_gpu.py tests/test_nbody_tri_gpu.py -x -q -k "every_variant or ragged" > gpurun_out/r3d_pytest.log 2>&1; echo "pytest_exit=$?"; tail -3 gpurun_out/r3d_pytest.log
timeout 400 python tools/nbody_crow_sweep.py 65536,131072 2022:16384:8:32 3002:16384:8:32 3022:16384:8:32 3022:4096:4:16 3032:16384:8:32 > gpurun_out/r3d_crow_sweep_small.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r3d_crow_sweep_small.txt
timeout 400 python tools/nbody_crow_sweep.py 1048576 2022:16384:8:32 2022:16384:4:32 2021:16384:8:32 3002:4096:2:8 3022:16384:8:32 3022:8192:4:16 3022:4096:2:8 3032:8192:4:16 3021:8192:4:16 > gpurun_out/r3d_crow_sweep_1m.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r3d_crow_sweep_1m.txt
