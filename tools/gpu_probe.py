"""Development probe run on the GPU box: quick parity checks + kernel variant sweeps.
Not part of the product or the test-suite; writes gpurun_out/probe.json."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402
from oracle import port  # noqa: E402

out = {}
L = H.lib()


def section(name):
    print(f"== {name}", flush=True)


section("device")
print("devices", L.hgb_device_count(), L.hgb_version())
tf, mhz = C.c_double(), C.c_double()
H._abi.check(L.hgb_measure_fp64_peak(C.byref(tf), C.byref(mhz)))
gbs = C.c_double()
H._abi.check(L.hgb_measure_hbm_copy(C.byref(gbs)))
out["fp64_tflops"], out["fp64_mhz_est"], out["hbm_copy_gbs"] = tf.value, mhz.value, gbs.value
print(out)

section("inv_sqrt selftest")
mm, fb = C.c_ulonglong(), C.c_double()
t0 = time.time()
H._abi.check(L.hgb_selftest_inv_sqrt(40_000_000_000, 12345, C.byref(mm), C.byref(fb)))
out["selftest_mismatches"] = mm.value
print("mismatches", mm.value, "first_bad", fb.value.hex(), "secs", time.time() - t0)

section("nbody parity")
ok = True
for n, it in [(3, 10), (3, 3), (1, 2), (100, 4), (1000, 5), (4097, 3), (5000, 2)]:
    p0 = synth.fill_planets(n) if n == 3 else synth.uniform_cube(n, seed=n)
    a, b = p0.copy(), np.zeros_like(p0)
    port.nbody_run(a, b, it)
    for blk, unr in [(0, 0), (64, 1), (256, 4), (128, 4)]:
        ga, gb = p0.copy(), np.zeros_like(p0)
        with H.nbody.NBody(n) as nb:
            nb.tune(blk, unr).upload(ga)
            nb.step(it)
            cur, prev = nb.download(want_prev=True)
        want_cur, want_prev = (a, b) if it % 2 == 0 else (b, a)
        same = np.array_equal(cur, want_cur) and (it < 1 or np.array_equal(prev, want_prev))
        ok &= same
        if not same:
            print("MISMATCH nbody", n, it, blk, unr, np.abs(cur - want_cur).max())
    ga, gb = p0.copy(), np.zeros_like(p0)
    H.nbody.run(ga, gb, n, it, 1)
    same = np.array_equal(ga, a) and np.array_equal(gb, b)
    ok &= same
    if not same:
        print("MISMATCH nbody drop-in", n, it)
out["nbody_parity"] = bool(ok)
print("nbody parity", ok)

section("conv parity")
ok = True
cases = [(3, 3, 1, synth.LAPLACE_TXT, 15), (17, 9, 2, synth.default_kernel(), 3),
         (33, 20, 3, synth.kernel_dense(7), 4), (8, 5, 2, synth.kernel_dense(4), 2),
         (300, 70, 1, synth.kernel_dense(3), 3), (513, 260, 2, synth.kernel_dense(5), 2),
         (700, 131, 3, synth.kernel_dense(7), 2), (256, 128, 1, synth.kernel_dense(3), 1),
         (512, 512, 2, synth.KERNEL_TXT, 10)]
for (w, h, p, k, it) in cases:
    img = synth.image_float(w, h, seed=w)
    want_c = synth.add_padding(img, p)
    want_p = np.full_like(want_c, 7.0)
    port.conv_run(want_c, want_p, w, h, p, k, it)
    for variant in (0, 1):
        c = synth.add_padding(img, p)
        with H.convolution.Conv2D(w, h, p) as cv:
            cv.set_kernel(k).tune(variant, 0).upload(c)
            cv.step(it)
            gc, gp = cv.download(want_prev=True)
        same = np.array_equal(gc, want_c) and np.array_equal(gp, want_p)
        ok &= same
        if not same:
            bad = np.argwhere(gc != want_c)
            print("MISMATCH conv", w, h, p, it, "variant", variant, len(bad), bad[:4].tolist())
out["conv_parity"] = bool(ok)
print("conv parity", ok)

section("nbody sweep N=65536")
n = 65536
pos = synth.plummer(n)
res = {}
for mode in (H.NBODY_EXACT, H.NBODY_FAST):
    for blk in (64, 128, 256):
        for unr in (1, 2, 4):
            with H.nbody.NBody(n, mode=mode) as nb:
                nb.tune(blk, unr).upload(pos)
                nb.step(1)
                ms = min(nb.step(1) for _ in range(3))
            rate = n * n / (ms * 1e-3)
            res[f"mode{mode}_b{blk}_u{unr}"] = rate
            print(f"mode={mode} block={blk} unroll={unr}: {ms:.3f} ms  {rate:.4e} inter/s", flush=True)
out["nbody_sweep_64k"] = res

section("nbody N=262144 best-ish")
n = 262144
pos = synth.uniform_cube(n)
for mode in (H.NBODY_EXACT, H.NBODY_FAST):
    for blk, unr in [(128, 2), (128, 4), (256, 2), (64, 4)]:
        with H.nbody.NBody(n, mode=mode) as nb:
            nb.tune(blk, unr).upload(pos)
            nb.step(1)
            ms = nb.step(2) / 2
        print(f"mode={mode} block={blk} unroll={unr}: {ms:.3f} ms  {n*n/(ms*1e-3):.4e} inter/s", flush=True)
        out[f"nbody_256k_mode{mode}_b{blk}_u{unr}"] = n * n / (ms * 1e-3)

section("conv sweep 8192x8192")
w = h = 8192
img = synth.image_float(w, h)
for K in (3, 5, 7):
    p = K // 2
    k = synth.kernel_dense(K)
    c = synth.add_padding(img, p)
    for variant, chunk in [(0, 64), (0, 128), (0, 256), (0, 512), (1, 0)]:
        with H.convolution.Conv2D(w, h, p) as cv:
            cv.set_kernel(k).tune(variant, chunk).upload(c)
            cv.step(2)
            ms = cv.step(5) / 5
        gbs_ = w * h * 16 / (ms * 1e-3) / 1e9
        print(f"K={K} variant={variant} chunk={chunk}: {ms:.3f} ms  {gbs_:.1f} GB/s algorithmic", flush=True)
        out[f"conv8k_K{K}_v{variant}_c{chunk}"] = gbs_

out["launches"] = L.hgb_launch_count()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/probe.json", "w"), indent=1)
print("DONE")
