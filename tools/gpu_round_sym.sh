#!/bin/bash
# GPU-box visit for the pair-symmetric nbody kernel: its parity tests (bounded by timeout: the kernel
# spins on flags, a bug must not hang the box), the variant sweep, then the regular suite + bench.
mkdir -p gpurun_out
HGB_TEST_SYM=1 timeout 200 python -m pytest tests/test_nbody_sym_gpu.py -q -x > gpurun_out/pytest_sym.log 2>&1; echo "sym_pytest_exit=$?"
tail -5 gpurun_out/pytest_sym.log
timeout 240 python tools/nbody_sym_sweep.py > gpurun_out/sym_sweep.txt 2>&1; echo "sweep_exit=$?"
cat gpurun_out/sym_sweep.txt
if [ "${FULL:-1}" = 1 ]; then
  timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_exit=$?"
  tail -3 gpurun_out/pytest_gpu.log
  timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench_exit=$?"
  HGB_NBODY_SYM=1 timeout 400 python bench.py --skip-conv --skip-cpu > gpurun_out/bench_sym.json 2> gpurun_out/bench_sym.err; echo "bench_sym_exit=$?"
fi
ls -la gpurun_out | tail -12
