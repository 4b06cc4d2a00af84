"""Small invocations of every kernel family for compute-sanitizer (one tool per run):
   compute-sanitizer --tool racecheck|memcheck python tools/sanitize_target.py
Each case is also checked against the oracle, so a run that exits 0 computed the right bits."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402
from oracle import port  # noqa: E402

n = 4097
pos = synth.plummer(n, seed=3)
want = np.zeros_like(pos)
port.nbody_step_range(pos, want, 0, n)
for name, variant, seg in (("nbody_step_kernel", 0, 0), ("nbody_sym_kernel<2,4,2>", 242, 0), ("nbody_tri_kernel<2>", 2002, 256)):
    with H.nbody.NBody(n) as nb:
        nb.symmetric(variant, 0)
        if seg:
            nb.segment(seg, 4, 2)
        nb.upload(pos)
        nb.step(1)
        ok = np.array_equal(nb.download(), want)
        print(f"{name}: n={n} one step, bit-identical to the oracle: {ok}  [{nb.describe()[0]}]", flush=True)
        assert ok
w, h = 300, 70
for k, mode, iters in ((3, H.CONV_EXACT, 4), (7, H.CONV_EXACT, 3), (5, H.CONV_EXACT, 3), (3, H.CONV_FAST, 4), (7, H.CONV_FAST, 3)):
    p = k // 2
    kern = synth.kernel_dense(k)
    img = synth.add_padding(synth.image_float(w, h, seed=w + h), p)
    a, b = img.copy(), np.zeros_like(img)
    port.conv_run(a, b, w, h, p, kern, iters)
    with H.convolution.Conv2D(w, h, p) as cv:
        cv.set_kernel(kern)
        cv.set_mode(mode)
        cv.upload(img)
        cv.step(iters)
        got = cv.download()
    if mode == H.CONV_EXACT:
        ok = np.array_equal(got, a)
    else:
        ok = np.abs(got - a).max() / np.abs(a).max() <= 1e-12
    print(f"conv {k}x{k} {'EXACT' if mode == H.CONV_EXACT else 'FAST'}: {w}x{h}, {iters} iterations, matches the oracle: {ok}", flush=True)
    assert ok
H.lib().hgb_finalize()
print("SANITIZE_TARGET_OK", flush=True)
