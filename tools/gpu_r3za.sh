#!/bin/bash
mkdir -p gpurun_out
LIGHT="--section WarpStateStats --section SchedulerStats --section ComputeWorkloadAnalysis --section SpeedOfLight --section MemoryWorkloadAnalysis --section LaunchStats --section Occupancy --metrics dram__bytes_read.sum,dram__bytes_write.sum"
timeout 120 python tools/prof_target.py sym 6242 > gpurun_out/r3za_plain_seg.log 2>&1; echo "plain=$?"; cat gpurun_out/r3za_plain_seg.log
timeout 400 ncu $LIGHT --clock-control none -k regex:nbody_seg_kernel -c 1 -f -o gpurun_out/r3za_prof_seg6242 python tools/prof_target.py sym 6242 > gpurun_out/r3za_ncu_seg.log 2>&1; echo "ncu=$?"; tail -2 gpurun_out/r3za_ncu_seg.log
