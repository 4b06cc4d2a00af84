#!/bin/bash
mkdir -p gpurun_out
G=""
for seg in 8192 16384 32768; do for rows in 1 2 4 8; do for w in 32 128 512; do G="$G 3002:$seg:$rows:$w"; done; done; done
timeout 600 python tools/nbody_crow_sweep.py 1048576 $G 3012:16384:1:128 3012:16384:8:32 > gpurun_out/r3i_crow_schedule_1m.txt 2>&1; echo "rc=$?"; cat gpurun_out/r3i_crow_schedule_1m.txt
G=""
for seg in 4096 16384; do for rows in 1 2 4 8 16; do for w in 16 64 256; do G="$G 3002:$seg:$rows:$w"; done; done; done
timeout 300 python tools/nbody_crow_sweep.py 131072 $G > gpurun_out/r3i_crow_schedule_131k.txt 2>&1; echo "rc=$?"; cat gpurun_out/r3i_crow_schedule_131k.txt
