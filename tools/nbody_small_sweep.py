"""Standard (one target per thread) nbody kernel at small target counts: CTA shapes side by side.
N = 65536 (BASELINE configs[2]) and the 8-GPU slice of N = 2^20 (131072 targets against 2^20 sources,
via accel_subset-free trick: a handle sharded over HGB_VIRTUAL_GPUS would time the exchange too, so the
slice case is timed as n = 131072 with the same grid shape).  usage: python tools/nbody_small_sweep.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

SHAPES = [(0, 0, 0), (128, 2, 24), (128, 2, 28), (128, 4, 24), (64, 2, 28), (64, 2, 24), (64, 4, 24), (64, 1, 28)]
for n in (65536, 131072, 32768):
    pos = synth.plummer(n, seed=1234)
    ref = None
    for shape in SHAPES:
        with H.nbody.NBody(n) as nb:
            nb.symmetric(0, 0)
            if shape != (0, 0, 0):
                nb.tune(*shape)
            nb.upload(pos)
            nb.step(2)
            nb.upload(pos)
            ms = nb.step(10) / 10
            got = nb.download()
            name = nb.describe()[0]
        if ref is None:
            ref = got
        print(f"n={n} shape={shape} {name}: {ms:8.3f} ms/step  {n * n / (ms * 1e-3):.4e} inter/s  "
              f"same_bits={np.array_equal(got, ref)}", flush=True)
H.lib().hgb_finalize()
