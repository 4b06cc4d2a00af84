#!/bin/bash
mkdir -p gpurun_out
TRI_GRID=small timeout 900 python tools/nbody_tri_multi.py 1048576 1 8 4 > gpurun_out/r2m_tri_multi.txt 2>&1; echo "multi_exit=$?"; cat gpurun_out/r2m_tri_multi.txt
