#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/prof_target.py sym 3002 262144 > gpurun_out/r3e_plain_crow.log 2>&1; echo "plain=$?"; cat gpurun_out/r3e_plain_crow.log
timeout 600 ncu --set full --import-source on --clock-control none -k regex:nbody_crow_kernel -c 1 -f -o gpurun_out/r3e_prof_crow3002 python tools/prof_target.py sym 3002 262144 > gpurun_out/r3e_ncu_crow.log 2>&1; echo "ncu=$?"; tail -3 gpurun_out/r3e_ncu_crow.log
ls -la gpurun_out | grep r3e
