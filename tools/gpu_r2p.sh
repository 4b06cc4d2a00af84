#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/conv_stream_sweep.py > gpurun_out/r2p_conv_stream_sweep.txt 2>&1; echo "exit=$?"; cat gpurun_out/r2p_conv_stream_sweep.txt
