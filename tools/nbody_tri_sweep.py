"""Second-generation pair-symmetric nbody step (csrc/nbody_tri.cuh): device time per variant (CTAs per SM)
and segment size beside the standard kernel and the first-generation wavefront, same inputs, results
compared.  usage: python tools/nbody_tri_sweep.py [n ...]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

sizes = [int(a) for a in sys.argv[1:]] or [1 << 20, 1 << 18, 1 << 16]


def run(n, pos, variant, seg=0, steps=1, rows=0, weight=0):
    with H.nbody.NBody(n) as nb:
        nb.symmetric(variant, 0)
        if variant >= 2000:
            nb.segment(seg, rows, weight)
        nb.upload(pos)
        nb.step(1)
        nb.upload(pos)
        ms = nb.step(steps) / steps
        out = nb.download()
        counters = None
        if variant >= 2010:
            c = (C.c_ulonglong * 7)()
            H._abi.check(H.lib().hgb_nbody_sym_counters(nb._h, c))
            counters = list(c)
    return ms, out, counters


for n in sizes:
    pos = synth.uniform_cube(n, seed=1)
    steps = 1 if n >= (1 << 19) else 5
    ms, ref, _ = run(n, pos, 0, steps=steps)
    print(f"n={n} standard             : {ms:9.3f} ms  {n * n / (ms * 1e-3):.4e} inter/s", flush=True)
    ms, out, _ = run(n, pos, 242, steps=steps)
    print(f"n={n} sym 242              : {ms:9.3f} ms  {n * n / (ms * 1e-3):.4e} inter/s  same={np.array_equal(out, ref)}", flush=True)
    grid = [(2002, 8192, 16, 16), (2002, 8192, 8, 16), (2002, 16384, 16, 16), (2002, 16384, 8, 32), (2002, 32768, 8, 32),
            (2002, 16384, 4, 32), (2001, 16384, 8, 32), (2003, 16384, 8, 32)]
    for variant, seg, rows, weight in grid:
        if seg > n:
            continue
        ms, out, _ = run(n, pos, variant, seg, steps=steps, rows=rows, weight=weight)
        print(f"n={n} tri {variant} seg={seg:6d} rows={rows:3d} weight={weight:3d}: {ms:9.3f} ms  {n * n / (ms * 1e-3):.4e} inter/s  "
              f"same={np.array_equal(out, ref)}", flush=True)
    for variant in (2012, 2013):
        ms, out, c = run(n, pos, variant, 8192 if n >= 8192 else 0, steps=1)
        tot = max(1, c[2])
        print(f"n={n} tri {variant} (profiling): {ms:9.3f} ms; of warp time in items: chain-word waits {100.0 * c[0] / tot:.1f} %, "
              f"row-accumulator waits {100.0 * c[1] / tot:.1f} %, triangles {100.0 * c[3] / tot:.1f} %, items {c[4]}, "
              f"cycles per item {tot / max(1, c[4]):.0f}", flush=True)
H.lib().hgb_finalize()
