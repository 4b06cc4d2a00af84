#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_nbody_crow_gpu.py tests/test_nbody_tri_gpu.py -x -q -k "every_variant or ragged or dealt or out_of_range" > gpurun_out/r3v_pytest.log 2>&1; echo "pytest_exit=$?"; tail -2 gpurun_out/r3v_pytest.log
timeout 400 python tools/nbody_crow_sweep.py 131072,1048576 3002:0:0:0 2002:16384:2:128 2002:16384:8:32 > gpurun_out/r3v_lds_if.txt 2>&1; echo "rc=$?"; grep -v "standard\|sym 242" gpurun_out/r3v_lds_if.txt
