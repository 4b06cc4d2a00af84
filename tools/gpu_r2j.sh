#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2j_gpus.txt; nvidia-smi topo -m >> gpurun_out/r2j_gpus.txt 2>&1
timeout 900 python tools/nbody_tri_multi.py 1048576 1 8 4 > gpurun_out/r2j_tri_multi.txt 2>&1; echo "multi_exit=$?"; cat gpurun_out/r2j_tri_multi.txt
