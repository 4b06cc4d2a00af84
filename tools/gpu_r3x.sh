#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_nbody_seg_gpu.py -x -q -k "not full_size" > gpurun_out/r3x_pytest_seg.log 2>&1; echo "pytest_exit=$?"; tail -4 gpurun_out/r3x_pytest_seg.log
timeout 300 python tools/nbody_crow_sweep.py 524288,1048576 4242:0:0:0 6242:0:0:0 6242:32768:2:128 6242:16384:4:64 7242:0:0:0 > gpurun_out/r3x_seg_pipe.txt 2>&1; echo "rc=$?"; grep -v "standard\|tri 2002" gpurun_out/r3x_seg_pipe.txt
