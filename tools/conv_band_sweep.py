"""One 8-GPU band on one GPU: a 16384 x 2048 image (what each rank of an 8-way split of 16384^2 computes),
time per iteration against the rows per CTA, beside 1/8 of the full image's time.  No exchange here: this
isolates the small-band loss (waves, tails) from the halo exchange.  usage: python tools/conv_band_sweep.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

W = 16384
FINE = {3: (36, 40, 44, 48, 52, 56, 60), 5: (48, 64, 80, 96, 128), 7: (52, 56, 60, 64, 69, 74, 80, 86, 93, 103, 128, 147)}
fine = "fine" in sys.argv
for k in ((3, 5, 7) if fine else (3, 7)):
    p = k // 2
    kern = synth.kernel_dense(k)
    for rows in (16384, 2048):
        img = synth.add_padding(synth.image_float(W, rows, seed=3), p)
        for chunk in ((0,) if rows == 16384 else ((0,) + FINE[k]) if fine else (0, 16, 24, 32, 48, 64, 96, 114, 128, 171, 228, 256, 342, 512)):
            if k == 7 and chunk and chunk < 32:
                continue
            with H.convolution.Conv2D(W, rows, p) as cv:
                cv.set_kernel(kern)
                if chunk:
                    cv.tune(0, chunk)
                cv.upload(img)
                cv.step(20)
                ms = min(cv.step(100), cv.step(100)) / 100
            ideal = "" if rows == 16384 else f"  ({ms * 1e3 / (full / 8 * 1e3):.3f} x an eighth of the full image)"
            if rows == 16384:
                full = ms
            print(f"{k}x{k} {W}x{rows} rows/CTA {chunk if chunk else 'default':>7}: {ms * 1e3:8.2f} us/iteration{ideal}", flush=True)
        del img
H.lib().hgb_finalize()
