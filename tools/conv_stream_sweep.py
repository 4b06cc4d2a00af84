"""hgb_conv_run streamed in row chunks: end-to-end time of the call (host buffers in, host buffers out) against
the chunk size, 16384^2 image, 10 iterations.  usage: python tools/conv_stream_sweep.py [size]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

cn = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
it = 10
dense = synth.image_float(cn, cn, seed=3)
for k in (3, 7):
    p = k // 2
    kern = synth.kernel_dense(k)
    band_t = torch.empty((cn, cn + 2 * p), dtype=torch.float64).pin_memory()
    got_t = torch.empty((cn, cn + 2 * p), dtype=torch.float64).pin_memory()
    band, got = band_t.numpy(), got_t.numpy()
    band[...] = synth.add_padding(dense, p)
    ref = None
    for rows in (0, 160, 255, 384, 512, 1024):
        if rows and rows < 8 * it * p and rows != 255:
            pass
        H.convolution.set_streaming(rows)
        best = 1e9
        for rep in range(3):
            got[...] = band
            t0 = time.perf_counter()
            H.convolution.run_flat(got, None, cn, cn, p, kern, it, 1)
            best = min(best, (time.perf_counter() - t0) * 1e3)
        if ref is None:
            ref = got.copy()
        same = np.array_equal(got, ref)
        print(f"{k}x{k} chunk rows {rows:5d} ({'one piece' if rows == 0 else 'streamed'}): {best:8.2f} ms  "
              f"{cn * cn * 16 * it / (best * 1e-3) / 1e9:7.1f} GB/s end to end  same={same}", flush=True)
    del band_t, got_t
H.convolution.set_streaming(1)
H.lib().hgb_finalize()
