#!/bin/bash
# full GPU suite + smoke + bench on one GPU with the final default policy
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r3z_pytest_gpu.log 2>&1; echo "pytest_exit=$?"; tail -4 gpurun_out/r3z_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3z_smoke.log 2>&1; echo "smoke_exit=$?"; tail -1 gpurun_out/r3z_smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r3z_bench.json 2> gpurun_out/r3z_bench.err; echo "bench_exit=$?"; tail -3 gpurun_out/r3z_bench.err; cut -c1-700 gpurun_out/r3z_bench.json
