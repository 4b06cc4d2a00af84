"""Dev probe: per-iteration overhead of the multi-GPU convolution step when compute is tiny.
torchrun --nproc-per-node 2 tools/conv_latency.py   (rank mode)   or   python tools/conv_latency.py sp"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

sp = len(sys.argv) > 1 and sys.argv[1] == "sp"
W = 16384
for rows_total in (int(os.environ.get("HROWS", "1024")), int(os.environ.get("HROWS2", "4096"))):
    for k in (3,):
        p = k // 2
        if sp:
            g = H.lib().hgb_device_count()
            for exch in (H.EXCHANGE_NCCL, H.EXCHANGE_P2P):
                for variant in (0, 2):
                    cv = H.convolution.Conv2D(W, rows_total, p, ngpus=g)
                    cv.set_exchange(exch).set_kernel(synth.kernel_dense(k)).tune(variant, 0)
                    cv.upload(synth.add_padding(synth.image_float(W, rows_total), p))
                    cv.step(5)
                    t0 = time.perf_counter()
                    ms = cv.step(50)
                    wall = (time.perf_counter() - t0) * 1e3
                    print(f"SP g={g} H={rows_total} exch={exch} variant={variant}: dev {ms / 50 * 1e3:.1f} us/iter  wall {wall / 50 * 1e3:.1f} us/iter", flush=True)
                    cv.close()
        else:
            rank, world = H.dist.init_rank_comm()
            for variant in (0, 2):
                cv = H.convolution.Conv2D(W, rows_total, p, ngpus=None)
                r0, rows = cv.band()
                cv.set_kernel(synth.kernel_dense(k)).tune(variant, 0)
                cv.upload(synth.add_padding(synth.image_float(W, rows_total)[r0:r0 + rows], p))
                cv.step(5)
                t0 = time.perf_counter()
                ms = cv.step(50)
                wall = (time.perf_counter() - t0) * 1e3
                if rank == 0:
                    print(f"RANK world={world} H={rows_total} variant={variant}: dev {ms / 50 * 1e3:.1f} us/iter  wall {wall / 50 * 1e3:.1f} us/iter", flush=True)
                cv.close()
H.lib().hgb_finalize()
