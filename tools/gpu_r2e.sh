#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_nbody_tri_gpu.py -x -q > gpurun_out/r2e_pytest_tri.log 2>&1; echo "pytest_exit=$?"; tail -4 gpurun_out/r2e_pytest_tri.log
timeout 900 python tools/nbody_tri_sweep.py 1048576 65536 > gpurun_out/r2e_tri_sweep.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r2e_tri_sweep.txt
