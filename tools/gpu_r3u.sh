#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_nbody_sym_gpu.py -x -q > gpurun_out/r3u_pytest_sym.log 2>&1; echo "pytest_exit=$?"; tail -2 gpurun_out/r3u_pytest_sym.log
timeout 100 python tools/sym_hist.py 1048576 2>&1 | grep "raw\|n=\|ahead"
timeout 200 python tools/sym_hyst.py 1048576
timeout 100 python tools/sym_hyst.py 524288 | head -3
