#!/bin/bash
# One GPU-box visit: GPU tests, the bench line, then (after both exited 0) the ncu launch list of
# the same bench command and one --set full capture per hot kernel.  Outputs -> gpurun_out/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_exit=$?"
tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; be=$?; echo "bench_exit=$be"
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref_exit=$?"
if [ "$be" = 0 ] && [ "${SKIP_NCU:-0}" = 0 ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 1 --warmup 1 > gpurun_out/bench_under_ncu.log 2>&1
  # the N = 2^20 nbody launch takes ~14 min under --set full (2 s x ~45 passes): only on request
  if [ "${NB_NCU:-0}" = 1 ]; then
    python tools/prof_target.py nbody > gpurun_out/pt_nbody.log 2>&1 && \
    ncu --set full --clock-control none --import-source on -k regex:nbody_step -c 1 -f -o gpurun_out/prof_nbody \
        python tools/prof_target.py nbody > gpurun_out/ncu_nbody.log 2>&1
  fi
  python tools/prof_target.py conv > gpurun_out/pt_conv.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:conv -c 8 -f -o gpurun_out/prof_conv \
      python tools/prof_target.py conv > gpurun_out/ncu_conv.log 2>&1
fi
ls -la gpurun_out | tail -20
