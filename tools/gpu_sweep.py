"""Development sweep of nbody kernel variants on the GPU box (not part of the product)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402
from oracle import port  # noqa: E402

sizes = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "131072,1048576").split(",")]
modes = [H.NBODY_EXACT, H.NBODY_FAST] if "fast" in sys.argv else [H.NBODY_EXACT]

# parity first
pos = synth.plummer(3000, seed=7)
want = np.zeros_like(pos)
port.nbody_step_range(pos, want, 0, 3000)
for blk in (128, 256):
    for unr in (1, 2, 4):
        for w in (24, 32, 40):
            with H.nbody.NBody(3000) as nb:
                nb.tune(blk, unr, w).upload(pos)
                nb.step(1)
                assert np.array_equal(nb.download(), want), (blk, unr, w)
print("parity ok", flush=True)

for n in sizes:
    pos = synth.uniform_cube(n, seed=1)
    for mode in modes:
        for blk in (128, 256):
            for unr in (1, 2, 4):
                for w in (24, 32, 40):
                    with H.nbody.NBody(n, mode=mode) as nb:
                        nb.tune(blk, unr, w).upload(pos)
                        if n <= 262144:
                            nb.step(1)
                        ms = nb.step(1)
                    print(f"n={n} mode={mode} block={blk} unroll={unr} warps={w}: {ms:9.3f} ms  "
                          f"{n * n / (ms * 1e-3):.4e} inter/s", flush=True)
