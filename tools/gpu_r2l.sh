#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_nbody_tri_gpu.py -x -q -l -k "dealt_over_gpus" > gpurun_out/r2l_a.log 2>&1; echo "a_exit=$?"; grep -a "HgbError\|passed\|failed\|^g \|^rows \|^n \|^seg " gpurun_out/r2l_a.log | tail -12
