#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_nbody_seg_gpu.py -x -q -k "not full_size and not 64k" > gpurun_out/r3w_pytest_seg.log 2>&1; echo "pytest_exit=$?"; tail -4 gpurun_out/r3w_pytest_seg.log
timeout 300 python tools/nbody_crow_sweep.py 131072,1048576 4242:0:0:0 4242:16384:1:64 4242:8192:2:128 4242:32768:2:128 4242:16384:8:32 5242:0:0:0 > gpurun_out/r3w_seg_sweep.txt 2>&1; echo "rc=$?"; grep -v "standard\|tri 2002" gpurun_out/r3w_seg_sweep.txt
