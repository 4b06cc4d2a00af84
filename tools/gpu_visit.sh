#!/bin/bash
# Round-end GPU visit with the pair-symmetric kernel as the 1-GPU default: full GPU suite, bench line,
# light ncu capture of the symmetric kernel (a --set full replay of this 1.4 s spin-waiting kernel does not
# finish in minutes), ncu launch list of the nbody part of the bench.
mkdir -p gpurun_out
timeout 330 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_exit=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench_exit=$?"; cut -c1-400 gpurun_out/bench.json
timeout 90 ncu --section WarpStateStats --section SchedulerStats --section ComputeWorkloadAnalysis --section SpeedOfLight \
    --section MemoryWorkloadAnalysis --section LaunchStats --section Occupancy --clock-control none \
    -k regex:nbody_sym_kernel -c 1 -f -o gpurun_out/prof_sym_242 python tools/prof_target.py sym 242 > gpurun_out/ncu_sym_242.log 2>&1; echo "ncu_sym=$?"
timeout 80 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_nbody.csv \
    python bench.py --steps 1 --warmup 1 --skip-conv --skip-cpu > gpurun_out/bench_under_ncu.log 2>&1; echo "ncu_list=$?"
ls -la gpurun_out | tail -8
