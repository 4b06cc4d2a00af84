#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_nbody_sym_gpu.py -x -q > gpurun_out/r3s_pytest_sym.log 2>&1; echo "pytest_exit=$?"; tail -2 gpurun_out/r3s_pytest_sym.log
timeout 300 python tools/sym_hyst.py 1048576 > gpurun_out/r3s_sym_early_close.txt 2>&1; cat gpurun_out/r3s_sym_early_close.txt
timeout 100 python tools/sym_hyst.py 524288 > gpurun_out/r3s_sym_early_close_512k.txt 2>&1; cat gpurun_out/r3s_sym_early_close_512k.txt
