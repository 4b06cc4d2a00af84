"""Pair-symmetric nbody step (csrc/nbody_sym.cuh): device time of every variant at N = 2^20 (and at
smaller N, where the wavefront ramp shows) beside the standard kernel, same inputs, results compared.
usage: python tools/nbody_sym_sweep.py [n ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

VARIANTS = (242, 241, 421, 181, 422, 243)
sizes = [int(a) for a in sys.argv[1:]] or [1 << 20, 1 << 19, 1 << 18]
for n in sizes:
    pos = synth.uniform_cube(n, seed=1)
    with H.nbody.NBody(n) as nb:
        nb.symmetric(0, 0).upload(pos)
        nb.step(1)
        nb.upload(pos)
        ms = nb.step(1)
        ref = nb.download()
    print(f"n={n} standard        : {ms:9.3f} ms  {n * n / (ms * 1e-3):.4e} inter/s", flush=True)
    for v in VARIANTS if n >= (1 << 20) else VARIANTS[:3]:
        with H.nbody.NBody(n) as nb:
            nb.symmetric(v, 0).upload(pos)
            nb.step(1)
            nb.upload(pos)
            ms = nb.step(1)
            same = np.array_equal(nb.download(), ref)
        print(f"n={n} symmetric {v:3d}: {ms:9.3f} ms  {n * n / (ms * 1e-3):.4e} inter/s  bit-identical={same}",
              flush=True)
H.lib().hgb_finalize()
