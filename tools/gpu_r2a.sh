#!/bin/bash
# Round 2, first GPU visit: full GPU suite (new: symbol-level shim parity, rank-mode world 1, 64-thread
# CTA variants, handle lifetime), the bench line with parity_check and the new configs, the small-N sweep.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/r2a_smi.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest_gpu.log 2>&1; echo "pytest_exit=$?"; tail -5 gpurun_out/r2a_pytest_gpu.log
timeout 400 python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench_exit=$?"; cut -c1-600 gpurun_out/r2a_bench.json; tail -3 gpurun_out/r2a_bench.err
timeout 200 python tools/nbody_small_sweep.py > gpurun_out/r2a_nbody_small_sweep.txt 2>&1; echo "sweep_exit=$?"; cat gpurun_out/r2a_nbody_small_sweep.txt
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err; echo "ref_exit=$?"; cut -c1-900 gpurun_out/r2a_bench_ref.json
ls -la gpurun_out | tail -8
