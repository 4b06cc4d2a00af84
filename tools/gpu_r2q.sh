#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/conv_band_sweep.py > gpurun_out/r2q_conv_band_sweep.txt 2>&1; echo "exit=$?"; cat gpurun_out/r2q_conv_band_sweep.txt
