"""nbody: 6 vs 7 resident CTAs per SM (warps = 24 / 28) at the per-rank target counts of 8, 4, 2, 1 GPUs.
Targets are sharded, sources are not: a rank of g GPUs runs N/g targets against all N sources; the wave
structure depends on the target count only, so N x N at N = 2^20 / g shows the same tail."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

for n in (131072, 262144, 524288, 1048576):
    pos = synth.uniform_cube(n, seed=1)
    ref = None
    for unr, w in ((2, 24), (1, 28), (2, 28), (4, 28), (2, 32)):
        with H.nbody.NBody(n) as nb:
            nb.tune(128, unr, w).upload(pos)
            if n <= 262144:
                nb.step(1)
                nb.upload(pos)
            ms = nb.step(1)
            got = nb.download() if n <= 262144 else None
        if got is not None:
            ref = got if ref is None else ref
            assert np.array_equal(got, ref)
        print(f"n={n} unroll={unr} warps={w}: {ms:9.3f} ms  {n * n / (ms * 1e-3):.4e} inter/s", flush=True)
H.lib().hgb_finalize()
