#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/prof_target.py sym 242 786432 > gpurun_out/r3p_plain_sym.log 2>&1; echo "plain=$?"; cat gpurun_out/r3p_plain_sym.log
timeout 800 ncu --set full --import-source on --clock-control none -k regex:nbody_sym_kernel -c 1 -f -o gpurun_out/r3p_prof_sym242 python tools/prof_target.py sym 242 786432 > gpurun_out/r3p_ncu_sym.log 2>&1; echo "ncu=$?"; tail -2 gpurun_out/r3p_ncu_sym.log
ls -la gpurun_out | grep r3p
