"""First-generation wavefront: falling `hyst` chunks behind after catching up.  python tools/sym_hyst.py [n]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
pos = synth.uniform_cube(n, seed=1)
with H.nbody.NBody(n) as nb:
    nb.symmetric(0, 0).upload(pos)
    nb.step(1)
    ref = nb.download()
for variant in (242, 1242):
    for hyst in ((0, 1, 2, 4) if variant == 242 else (0, 1)):
        os.environ["HGB_NBODY_SYM_HYST_NOW"] = str(hyst)
        with H.nbody.NBody(n) as nb:
            nb.symmetric(variant, 0).upload(pos)
            nb.step(1)
            nb.upload(pos)
            ms = nb.step(1)
            same = np.array_equal(nb.download(), ref)
            extra = ""
            if variant == 1242:
                c = (C.c_ulonglong * 7)()
                H._abi.check(H.lib().hgb_nbody_sym_counters(nb._h, c))
                c = list(c)
                extra = f"  wait {100 * c[0] / c[5]:.1f} % fast loop {100 * c[3] / c[5]:.1f} %, fast-loop cycles per chunk {c[3] / c[6]:.0f}"
        print(f"n={n} sym {variant} hyst={hyst:2d}: {ms:9.3f} ms  same={same}{extra}", flush=True)
H.lib().hgb_finalize()
