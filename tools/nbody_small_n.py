"""Where the cooperative-row kernel (nbody_crow_kernel) overtakes the standard kernel on one GPU: device time
per step for a ladder of sizes, results compared bit for bit.  usage: python tools/nbody_small_n.py [n ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

sizes = [int(a) for a in sys.argv[1:]] or [8192, 16384, 24576, 32768, 49152, 65536, 98304, 196608, 262144, 393216, 524288]


def run(n, pos, variant, seg=0, rows=0, weight=0, steps=5):
    with H.nbody.NBody(n) as nb:
        nb.symmetric(variant, 0)
        if variant >= 2000:
            nb.segment(seg, rows, weight)
        nb.upload(pos)
        nb.step(2)
        nb.upload(pos)
        ms = nb.step(steps) / steps
        return ms, nb.download()


for n in sizes:
    pos = synth.plummer(n, seed=1234) if n == 65536 else synth.uniform_cube(n, seed=1)
    steps = 2 if n >= (1 << 18) else 6
    ms0, ref = run(n, pos, 0, steps=steps)
    line = f"n={n:7d} standard {ms0:8.3f} ms"
    for variant, seg, rows, weight in ((3002, 16384, 8, 32), (3002, 4096, 8, 32), (3002, 4096, 4, 16), (3002, 2048, 8, 32), (2002, 16384, 8, 32), (242, 0, 0, 0)):
        ms, out = run(n, pos, variant, min(seg, max(32, n // 32 * 32)) if seg else 0, rows, weight, steps=steps)
        line += f" | {variant}:{seg}:{rows}:{weight} {ms:8.3f} ({ms0 / ms:4.2f}x{'' if np.array_equal(out, ref) else ' DIFF'})"
    print(line, flush=True)
H.lib().hgb_finalize()
