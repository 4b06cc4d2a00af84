"""Multi-GPU convolution: per-iteration time against halo depth (iterations per exchange).
  torchrun --nproc-per-node G tools/conv_depth_sweep.py          one process per GPU (NCCL send/recv)
  python tools/conv_depth_sweep.py sp                            one process driving all GPUs
ROWS="16384,4096" picks the image heights (4096 rows over 2 GPUs = the 8-GPU band height)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

sp = len(sys.argv) > 1 and sys.argv[1] == "sp"
W = 16384
ITERS = 96
heights = [int(x) for x in os.environ.get("ROWS", "16384,4096").split(",")]
if not sp:
    rank, world = H.dist.init_rank_comm()
else:
    rank, world = 0, H.lib().hgb_device_count()
for rows_total in heights:
    dense = synth.image_float(W, rows_total, seed=3)
    for k in (3, 7):
        p = k // 2
        modes = [("nccl", H.EXCHANGE_NCCL), ("p2p", H.EXCHANGE_P2P)] if sp else [("nccl", H.EXCHANGE_NCCL)]
        for ename, exch in modes:
            for depth in (1, 2, 4, 8, 16, 32):
                cv = H.convolution.Conv2D(W, rows_total, p, ngpus=world if sp else None)
                if sp:
                    cv.set_exchange(exch)
                r0, rows = cv.band()
                eff = cv.set_depth(depth)
                if eff != depth:
                    cv.close()
                    continue
                cv.set_kernel(synth.kernel_dense(k)).upload_dense(dense[r0:r0 + rows])
                cv.step(16)
                t0 = time.perf_counter()
                ms = cv.step(ITERS)
                wall = (time.perf_counter() - t0) * 1e3
                if rank == 0:
                    gbs = W * rows_total * 16.0 * ITERS / (ms * 1e-3) / 1e9
                    print(f"{'SP' if sp else 'RANK'} g={world} H={rows_total} k={k} {ename} depth={depth}: "
                          f"dev {ms / ITERS * 1e3:7.1f} us/iter  wall {wall / ITERS * 1e3:7.1f} us/iter  {gbs:8.1f} GB/s",
                          flush=True)
                cv.close()
H.lib().hgb_finalize()
