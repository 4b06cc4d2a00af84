"""Third-generation pair-symmetric nbody step (csrc/nbody_crow.cuh): device time per variant and schedule
beside the standard kernel, the first and the second generation; results compared bit for bit.
usage: python tools/nbody_crow_sweep.py n[,n...] [variant:seg:rows:weight ...]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

sizes = [int(a) for a in (sys.argv[1] if len(sys.argv) > 1 else "131072,262144,1048576").split(",")]
grid = [tuple(int(x) for x in a.split(":")) for a in sys.argv[2:]] or [
    (3002, 16384, 8, 32), (3002, 8192, 8, 32), (3002, 4096, 8, 32), (3002, 4096, 4, 16), (3002, 2048, 4, 16),
    (3002, 16384, 2, 8), (3002, 8192, 2, 8), (3002, 4096, 2, 8), (3002, 4096, 1, 4), (3002, 2048, 1, 4),
    (3001, 8192, 4, 16), (3001, 4096, 2, 8), (3101, 8192, 4, 16), (3101, 4096, 2, 8), (3101, 2048, 1, 4)]


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
        if (2000 <= variant < 4000 and (variant // 10) % 10 in (1, 3)) or variant in (5242, 7242, 7241):
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
    ms, out, _ = run(n, pos, 2002, 16384 if n > 16384 else 0, steps=steps, rows=8, weight=32)
    print(f"n={n} tri 2002 seg=16384 rows=8 weight=32: {ms:9.3f} ms  {n * n / (ms * 1e-3):.4e} inter/s  same={np.array_equal(out, ref)}", flush=True)
    for spec in grid:
        variant, seg, rows, weight = spec[:4]
        flags = spec[4] if len(spec) > 4 else 0
        os.environ["HGB_NBODY_TRI_FLAGS"] = str(flags)      # read by the library at every launch
        os.environ["HGB_NBODY_SEG_HYST"] = str(flags)
        if seg > n:
            continue
        ms, out, c = run(n, pos, variant, seg, steps=steps, rows=rows, weight=weight)
        extra = ""
        if c and variant in (5242, 7242, 7241):
            tot = max(1, c[2])
            extra = (f"  of item time: chain-word waits {100.0 * c[0] / tot:.1f} % row-sum waits {100.0 * c[1] / tot:.1f} % predicate-free loop "
                     f"{100.0 * c[3] / tot:.1f} % predicated groups {100.0 * c[4] / tot:.1f} %; items {c[5]}, loop cycles per chunk {c[3] / max(1, c[6]):.0f}")
        elif c:
            tot = max(1, c[2])
            extra = (f"  waits: chain {100.0 * c[0] / tot:.1f} % racc {100.0 * c[1] / tot:.1f} % token {100.0 * c[5] / tot:.1f} % "
                     f"triangles {100.0 * c[3] / tot:.1f} %")
        print(f"n={n} {'seg ' if variant >= 4000 else 'crow' if variant >= 3000 else 'tri '} {variant} seg={seg:6d} rows={rows:3d} weight={weight:3d} flags={flags}: {ms:9.3f} ms  {n * n / (ms * 1e-3):.4e} inter/s  "
              f"same={np.array_equal(out, ref)}{extra}", flush=True)
H.lib().hgb_finalize()
