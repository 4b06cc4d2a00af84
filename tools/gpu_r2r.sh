#!/bin/bash
# Round-2 profiles on one GPU: ncu --set full of the convolution kernels, light captures (pipe / issue / DRAM bytes)
# of the three nbody kernels at N = 2^20, launch list of the bench command.  Every target runs once without ncu first.
mkdir -p gpurun_out
LIGHT="--section WarpStateStats --section SchedulerStats --section ComputeWorkloadAnalysis --section SpeedOfLight --section MemoryWorkloadAnalysis --section LaunchStats --section Occupancy --metrics dram__bytes_read.sum,dram__bytes_write.sum"
timeout 120 python tools/prof_target.py conv > gpurun_out/r2r_plain_conv.log 2>&1; echo "plain_conv=$?"; cat gpurun_out/r2r_plain_conv.log
timeout 600 ncu --set full --import-source on --clock-control none -k regex:conv -c 8 -f -o gpurun_out/r2r_prof_conv python tools/prof_target.py conv > gpurun_out/r2r_ncu_conv.log 2>&1; echo "ncu_conv=$?"
timeout 120 python tools/prof_target.py sym 242 > gpurun_out/r2r_plain_sym.log 2>&1; echo "plain_sym=$?"; cat gpurun_out/r2r_plain_sym.log
timeout 300 ncu $LIGHT --clock-control none -k regex:nbody_sym_kernel -c 1 -f -o gpurun_out/r2r_prof_sym242 python tools/prof_target.py sym 242 > gpurun_out/r2r_ncu_sym.log 2>&1; echo "ncu_sym=$?"
timeout 120 python tools/prof_target.py sym 2002 > gpurun_out/r2r_plain_tri.log 2>&1; echo "plain_tri=$?"; cat gpurun_out/r2r_plain_tri.log
timeout 300 ncu $LIGHT --clock-control none -k regex:nbody_tri_kernel -c 1 -f -o gpurun_out/r2r_prof_tri2002 python tools/prof_target.py sym 2002 > gpurun_out/r2r_ncu_tri.log 2>&1; echo "ncu_tri=$?"
timeout 120 python tools/prof_target.py sym 0 > gpurun_out/r2r_plain_std.log 2>&1; echo "plain_std=$?"; cat gpurun_out/r2r_plain_std.log
timeout 300 ncu $LIGHT --clock-control none -k regex:nbody_step_kernel -c 1 -f -o gpurun_out/r2r_prof_std python tools/prof_target.py sym 0 > gpurun_out/r2r_ncu_std.log 2>&1; echo "ncu_std=$?"
timeout 300 python bench.py --steps 1 --warmup 3 --skip-cpu --skip-extra > gpurun_out/r2r_bench_plain.json 2> gpurun_out/r2r_bench_plain.err; echo "bench_plain=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2r_launches_bench.csv python bench.py --steps 1 --warmup 3 --skip-cpu --skip-extra > gpurun_out/r2r_bench_under_ncu.log 2>&1; echo "ncu_list=$?"
ls -la gpurun_out | grep r2r
