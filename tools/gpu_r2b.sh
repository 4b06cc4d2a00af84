#!/bin/bash
# Round 2, second GPU visit: the second-generation pair-symmetric kernel: parity tests first (small, with the
# in-kernel spin limit as the safety net), then the sweep.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_nbody_tri_gpu.py -x -q > gpurun_out/r2b_pytest_tri.log 2>&1; echo "pytest_exit=$?"; tail -15 gpurun_out/r2b_pytest_tri.log
timeout 900 python tools/nbody_tri_sweep.py 1048576 65536 > gpurun_out/r2b_tri_sweep.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r2b_tri_sweep.txt
