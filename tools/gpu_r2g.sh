#!/bin/bash
# ncu --set full of the second-generation pair-symmetric kernel at N = 2^19, source page included.
# The program is run once without ncu first.
mkdir -p gpurun_out
timeout 120 python tools/prof_target.py sym 2002 524288 > gpurun_out/r2g_plain.log 2>&1; echo "plain_exit=$?"; cat gpurun_out/r2g_plain.log
timeout 800 ncu --set full --import-source on --clock-control none -k regex:nbody_tri_kernel -c 1 -f -o gpurun_out/r2g_tri_2002_512k \
    python tools/prof_target.py sym 2002 524288 > gpurun_out/r2g_ncu.log 2>&1; echo "ncu_exit=$?"; tail -5 gpurun_out/r2g_ncu.log
ls -la gpurun_out/*.ncu-rep
