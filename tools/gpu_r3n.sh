#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_nbody_crow_gpu.py tests/test_nbody_tri_gpu.py -x -q -k "every_variant or ragged or dealt" > gpurun_out/r3n_pytest.log 2>&1; echo "pytest_exit=$?"; tail -2 gpurun_out/r3n_pytest.log
timeout 400 python tools/nbody_crow_sweep.py 131072,1048576 3002:0:0:0 3002:16384:8:32 2002:16384:2:128 3012:0:0:0 > gpurun_out/r3n_publish_early.txt 2>&1; echo "rc=$?"; cat gpurun_out/r3n_publish_early.txt
