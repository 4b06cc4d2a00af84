"""Short profiling target for ncu --set full: one nbody step (EXACT and FAST) and a few
convolution iterations.  Usage: python tools/prof_target.py [nbody|sym|conv]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "nbody"
if what == "nbody":
    # the bench workload itself (N = 2^20, EXACT): 8192 CTAs = 9.2 waves, so the tail is the bench's
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 1048576
    pos = synth.uniform_cube(n, seed=1)
    with H.nbody.NBody(n, mode=H.NBODY_EXACT) as nb:
        nb.upload(pos)
        print("EXACT ms", nb.step(1))
elif what == "sym":
    # pair-symmetric kernel: python tools/prof_target.py sym <variant> [n]
    variant = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 1048576
    pos = synth.uniform_cube(n, seed=1)
    with H.nbody.NBody(n, mode=H.NBODY_EXACT) as nb:
        nb.symmetric(variant, 0).upload(pos)
        print("SYM", variant, "ms", nb.step(1))
else:
    n = 16384
    dense = synth.image_float(n, n, seed=3)
    for k in (3, 7):
        p = k // 2
        with H.convolution.Conv2D(n, n, p) as cv:
            cv.set_kernel(synth.kernel_dense(k)).upload(synth.add_padding(dense, p))
            print("k", k, "ms", cv.step(5))     # 3x3: two fused passes (2 iterations each) + one plain pass
H.lib().hgb_finalize()
