#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_nbody_crow_gpu.py tests/test_nbody_tri_gpu.py tests/test_nbody_sym_gpu.py -x -q > gpurun_out/r3j_pytest.log 2>&1; echo "pytest_exit=$?"; tail -3 gpurun_out/r3j_pytest.log
timeout 400 python tools/nbody_crow_sweep.py 1048576 2002:16384:1:32 2002:16384:1:128 2002:16384:2:32 2002:16384:2:128 2002:8192:1:32 2002:32768:1:128 3002:0:0:0 > gpurun_out/r3j_tri_rows_1m.txt 2>&1; echo "rc=$?"; cat gpurun_out/r3j_tri_rows_1m.txt
timeout 300 python tools/nbody_crow_sweep.py 524288,262144 2002:16384:1:32 2002:16384:2:64 2002:16384:8:32 3002:0:0:0 3002:8192:1:32 > gpurun_out/r3j_tri_rows_mid.txt 2>&1; echo "rc=$?"; cat gpurun_out/r3j_tri_rows_mid.txt
