#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_nbody_seg_gpu.py tests/test_nbody_tri_gpu.py tests/test_nbody_crow_gpu.py -x -q -k "not full_size and not 64k and not epoch_wrap" > gpurun_out/r4a_pytest.log 2>&1; echo "pytest_exit=$?"; tail -5 gpurun_out/r4a_pytest.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29541 tools/rank_mode_check.py > gpurun_out/r4a_rank1.out 2> gpurun_out/r4a_rank1.err; echo "rank1_exit=$?"; grep "RANK_MODE_OK\|cases ran" gpurun_out/r4a_rank1.out; tail -3 gpurun_out/r4a_rank1.err
