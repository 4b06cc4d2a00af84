#!/bin/bash
# ncu captures of the pair-symmetric nbody kernel at N = 2^20: one --set full, two light ones.
mkdir -p gpurun_out
V=${V:-234}
python tools/prof_target.py sym $V > gpurun_out/pt_sym.log 2>&1; echo "plain_exit=$?"; cat gpurun_out/pt_sym.log | tail -2
timeout 400 ncu --set full --clock-control none --import-source on -k regex:nbody_sym_kernel -c 1 -f -o gpurun_out/prof_sym_$V \
    python tools/prof_target.py sym $V > gpurun_out/ncu_sym.log 2>&1; echo "ncu_exit=$?"
for W in ${LIGHT:-226 124}; do
  timeout 200 ncu --section WarpStateStats --section SchedulerStats --section ComputeWorkloadAnalysis --section SpeedOfLight \
      --clock-control none -k regex:nbody_sym_kernel -c 1 -f -o gpurun_out/prof_sym_light_$W \
      python tools/prof_target.py sym $W > gpurun_out/ncu_sym_$W.log 2>&1; echo "ncu_light_$W=$?"
done
ls -la gpurun_out | tail
