"""Pins the oracle (our CPU restatement) BEFORE it is trusted as the parity checker:
against the reference's own gtest golden matrices, against results of the unmodified
reference committed in tests/golden/golden.json, against oracle/_ref when it is present,
and against an independent numpy restatement.  CPU only."""
import hashlib
import os

import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth
from oracle import npref, port, ref

F = float.fromhex


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def unhex(xs, shape):
    return np.array([F(x) for x in xs], dtype=np.float64).reshape(shape)


# ---- nbody ---------------------------------------------------------------------------
@pytest.mark.parametrize("n,iters", [(3, 0), (3, 1), (3, 2), (3, 3), (3, 10), (5, 10), (16, 7)])
def test_nbody_fill_planet_golden(golden, n, iters):
    g = golden["nbody"][f"fill_planet_n{n}_it{iters}"]
    a, b = port.fill_planets(n), np.zeros((n, 3))
    port.nbody_run(a, b, iters, threads=2)
    assert np.array_equal(a, unhex(g["planets"], (n, 3)))
    assert np.array_equal(b, unhex(g["buffer"], (n, 3)))


def test_nbody_shipped_default_matches_printed_result(golden):
    # reference CLI defaults: 3 bodies, 10 iterations (src/nbody/nbody-util.c:27-28)
    a, b = port.fill_planets(3), np.zeros((3, 3))
    port.nbody_run(a, b, 10)
    lines = ["(%g, %g, %g)" % tuple(r) for r in a]
    assert lines == golden["cli"]["nbody -s 3 -n 10"]
    assert lines == ["(-0.0617896, -0.0123579, -1.85369)", "(0, 0, 0)", "(0.0617896, 0.0123579, 1.85369)"]


def test_nbody_odd_iterations_leave_previous_state_in_planets(golden):
    # the reference swaps LOCAL pointers: -n 2 and -n 3 print the same file
    assert golden["cli"]["nbody -s 3 -n 2"] == golden["cli"]["nbody -s 3 -n 3"]
    a2, b2 = port.fill_planets(3), np.zeros((3, 3))
    port.nbody_run(a2, b2, 2)
    a3, b3 = port.fill_planets(3), np.zeros((3, 3))
    port.nbody_run(a3, b3, 3)
    assert np.array_equal(a2, a3) and not np.array_equal(b2, b3)


@pytest.mark.parametrize("name,gen,n,iters", [("uniform", synth.uniform_cube, 257, 4),
                                              ("plummer", synth.plummer, 1000, 3),
                                              ("uniform", synth.uniform_cube, 2049, 10),
                                              ("plummer", synth.plummer, 4096, 10)])
def test_nbody_synthetic_golden(golden, name, gen, n, iters):
    g = golden["nbody"][f"{name}_n{n}_it{iters}"]
    a, b = gen(n), np.zeros((n, 3))
    port.nbody_run(a, b, iters)
    assert sha(a) == g["planets_sha256"] and sha(b) == g["buffer_sha256"]
    assert np.array_equal(a[:2], unhex(g["planets_head"], (2, 3)))


def test_nbody_thread_count_and_range_split_do_not_change_bits():
    pos = synth.plummer(700)
    full = np.zeros_like(pos)
    port.nbody_step_range(pos, full, 0, 700, threads=1)
    parts = np.zeros_like(pos)
    port.nbody_step_range(pos, parts, 0, 350, threads=3)
    port.nbody_step_range(pos, parts, 350, 700, threads=5)
    assert np.array_equal(full, parts)
    sub = port.nbody_accel_subset(pos, [0, 699, 13, 13])
    assert np.array_equal(sub, full[[0, 699, 13, 13]])


def test_nbody_numpy_restatement_agrees():
    for pos, iters in ((synth.fill_planets(3), 10), (synth.uniform_cube(130), 3)):
        a, b = pos.copy(), np.zeros_like(pos)
        port.nbody_run(a, b, iters)
        c, d = pos.copy(), np.zeros_like(pos)
        npref.nbody_run(c, d, iters)
        assert np.array_equal(a, c) and np.array_equal(b, d)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_nbody_port_equals_reference_binary():
    for pos, iters, procs in ((synth.fill_planets(7), 5, 1), (synth.plummer(900), 4, 4),
                              (synth.uniform_cube(1025), 3, 3)):
        a, b = pos.copy(), np.zeros_like(pos)
        port.nbody_run(a, b, iters)
        c, d = pos.copy(), np.zeros_like(pos)
        ref.nbody_run(c, d, iters, procs)
        assert np.array_equal(a, c) and np.array_equal(b, d)


def test_nbody_energy_extension_is_finite_and_negative():
    e = port.nbody_energy(synth.plummer(300))
    assert np.isfinite(e) and e < 0


# ---- convolution: the reference's own gtest vectors ---------------------------------------
IMG3 = np.array([[-1.5, 0, -1], [9, -2, 3], [2, 0, 5]], dtype=np.float64)
PADDED7 = np.array([[-1.5, -1.5, -1.5, 0, -1, -1, -1],
                    [-1.5, -1.5, -1.5, 0, -1, -1, -1],
                    [-1.5, -1.5, -1.5, 0, -1, -1, -1],
                    [9, 9, 9, -2, 3, 3, 3],
                    [2, 2, 2, 0, 5, 5, 5],
                    [2, 2, 2, 0, 5, 5, 5],
                    [2, 2, 2, 0, 5, 5, 5]], dtype=np.float64)


def test_add_padding_complicated_image():
    # test/convolutionUtilTests.cpp:96-134 -- rows 2..4 of the 7x7 golden are the stored rows
    flat = port.add_padding(IMG3, 2)
    assert np.array_equal(flat, PADDED7[2:5])
    assert np.array_equal(synth.add_padding(IMG3, 2), PADDED7[2:5])


def test_run_laplace_neutralizes_image(golden):
    # test/convolutionTests.cpp:9-64: 15 iterations of the 3x3 cross, checksum 14.5 -> 0
    assert port.conv_checksum(port.add_padding(IMG3, 1), 3, 3, 1) == 14.5
    cur = port.add_padding(IMG3, 1)
    prev = cur.copy()
    port.conv_run(cur, prev, 3, 3, 1, synth.LAPLACE_TXT, 15)
    assert port.conv_checksum(cur, 3, 3, 1) == 0.0
    g = golden["conv"]["gtest_laplace_15"]
    assert np.array_equal(cur, unhex(g["cur"], (3, 5))) and np.array_equal(prev, unhex(g["prev"], (3, 5)))


def test_update_borders_semantics():
    # test/convolutionUtilTests.cpp:184-231: after editing inner pixels, one pass of a
    # 1-tap identity kernel rewrites the pad columns from the edge pixels
    rows = PADDED7[2:5].copy()
    rows[0, 2:5] = [2, 0, 100]
    rows[1, 2:4] = [1, 0]
    rows[2, 3:5] = [-1, 5]
    ident = np.zeros((5, 5))
    ident[2, 2] = 1.0
    cur, prev = rows.copy(), np.zeros_like(rows)
    port.conv_run(cur, prev, 3, 3, 2, ident, 1)
    want = np.array([[2, 2, 2, 0, 100, 100, 100], [1, 1, 1, 0, 3, 3, 3], [2, 2, 2, -1, 5, 5, 5]], dtype=float)
    assert np.array_equal(cur, want)
    assert np.array_equal(prev, rows)          # one iteration: buffer holds state 0


# ---- convolution: results of the unmodified reference -------------------------------------
def _conv_case(golden, key, dense, kernel):
    g = golden["conv"][key]
    p, it = g["p"], g["iters"]
    h, w = dense.shape
    cur = port.add_padding(dense, p)
    prev = np.zeros_like(cur)
    port.conv_run(cur, prev, w, h, p, kernel, it)
    assert sha(cur) == g["cur_sha256"], key
    assert sha(prev) == g["prev_sha256"], key
    cs = port.conv_checksum(cur, w, h, p)
    assert cs == F(g["checksum"]) and "%f" % cs == g["checksum_printed"]


@pytest.mark.parametrize("it", [1, 10, 100])
def test_conv_shipped_image_kernel_txt(golden, shipped_image, it):
    _conv_case(golden, f"input_kernel_txt_n{it}", shipped_image, synth.KERNEL_TXT)
    assert golden["conv"][f"input_kernel_txt_n{it}"]["checksum_printed"] == "34779110.000000"


@pytest.mark.parametrize("it", [1, 10, 100])
def test_conv_shipped_image_laplace_txt(golden, shipped_image, it):
    _conv_case(golden, f"input_laplace_txt_n{it}", shipped_image, synth.LAPLACE_TXT)


def test_conv_shipped_image_default_kernel(golden, shipped_image):
    _conv_case(golden, "input_default5x5_n10", shipped_image, synth.default_kernel())
    assert golden["conv"]["input_default5x5_n10"]["checksum_printed"] == "-167053284437852.000000"


def test_conv_shipped_float_image(golden, shipped_image):
    _conv_case(golden, "txt_kernel_txt_n10", shipped_image / 256.0, synth.KERNEL_TXT)
    assert golden["conv"]["txt_kernel_txt_n10"]["checksum_printed"] == "135855.898438"


@pytest.mark.parametrize("w,h,k,it", [(33, 20, 7, 4), (300, 70, 3, 3), (513, 260, 5, 2),
                                      (1024, 768, 3, 5), (640, 480, 7, 3)])
def test_conv_synthetic_float_golden(golden, w, h, k, it):
    _conv_case(golden, f"synth_float_{w}x{h}_k{k}_n{it}", synth.image_float(w, h), synth.kernel_dense(k))


@pytest.mark.parametrize("k,it", [(3, 4), (7, 2)])
def test_conv_synthetic_int_golden(golden, k, it):
    _conv_case(golden, f"synth_int_512x384_k{k}_n{it}", synth.image_int(512, 384), synth.kernel_dense(k))


def test_conv_numpy_restatement_agrees():
    for (w, h, p, k, it) in [(17, 9, 2, synth.default_kernel(), 3), (40, 11, 3, synth.kernel_dense(7), 2),
                             (8, 5, 2, synth.kernel_dense(4), 2), (5, 1, 1, synth.kernel_dense(3), 3)]:
        base = synth.add_padding(synth.image_float(w, h), p)
        a, b = base.copy(), np.zeros_like(base)
        port.conv_run(a, b, w, h, p, k, it)
        c, d = base.copy(), np.zeros_like(base)
        npref.conv_run(c, d, w, h, p, k, it)
        assert np.array_equal(a, c) and np.array_equal(b, d)


def test_conv_row_split_does_not_change_bits():
    w, h, p = 50, 37, 2
    k = synth.kernel_dense(5)
    inp = synth.add_padding(synth.image_float(w, h), p)
    full, parts = np.zeros_like(inp), np.zeros_like(inp)
    port.conv_step_rows(inp, full, w, h, p, k, 0, h, threads=1)
    port.conv_step_rows(inp, parts, w, h, p, k, 0, 19, threads=2)
    port.conv_step_rows(inp, parts, w, h, p, k, 19, h, threads=3)
    assert np.array_equal(full, parts)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_conv_port_equals_reference_binary():
    for (w, h, p, k, it, procs) in [(3, 3, 1, synth.LAPLACE_TXT, 15, 1), (64, 48, 3, synth.kernel_dense(7), 3, 4),
                                    (10, 4, 2, synth.kernel_dense(4), 2, 2), (129, 33, 1, synth.kernel_dense(3), 1, 3),
                                    (31, 17, 2, synth.kernel_dense(5)[:3, :], 2, 1)]:
        img = synth.image_float(w, h, seed=w + h)
        a, b = port.add_padding(img, p), np.full((h, w + 2 * p), 3.0)
        port.conv_run(a, b, w, h, p, k, it)
        c, d = ref.add_padding(img, p), np.full((h, w + 2 * p), 3.0)
        ref.conv_run(c, d, w, h, p, k, it, procs)
        assert np.array_equal(a, c) and np.array_equal(b, d)
        assert port.conv_checksum(a, w, h, p) == ref.conv_checksum(c, w, h, p)


@pytest.mark.skipif(not (ref.available() and ref.available("_pgo")),
                    reason="oracle/_ref PGO build not present (make -C oracle ref-pgo)")
def test_pgo_build_of_the_reference_is_bit_identical_to_the_shipped_flags():
    """the -O3 + profile-feedback build (what pgo.sh intended) is a second CPU baseline for
    bench.py; it must compute exactly what the shipped-flags build computes"""
    try:
        res = {}
        for v in ("", "_pgo"):
            ref.use(v)
            out = []
            for pos, iters, procs in ((synth.fill_planets(9), 4, 1), (synth.plummer(1500), 3, 4)):
                a, b = pos.copy(), np.zeros_like(pos)
                ref.nbody_run(a, b, iters, procs)
                out += [a, b]
            for (w, h, k, it) in [(130, 70, 3, 3), (64, 48, 5, 2), (97, 33, 7, 2)]:
                p = k // 2
                c, d = ref.add_padding(synth.image_float(w, h, seed=k), p), np.zeros((h, w + 2 * p))
                ref.conv_run(c, d, w, h, p, synth.kernel_dense(k), it, 3)
                out += [c, d, np.array([ref.conv_checksum(c, w, h, p)])]
            res[v] = out
        for x, y in zip(res[""], res["_pgo"]):
            assert np.array_equal(x.view(np.uint64), y.view(np.uint64))
    finally:
        ref.use("")


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="reference tree not present")
def test_golden_fixture_is_the_shipped_file(shipped_image):
    img = ref.read_image("/root/reference/2d-convolution.input")
    assert np.array_equal(img, shipped_image)


# ---- integrator extension: pinned to the reference through the unit-mass case ----
def test_integrator_unit_mass_kick_is_the_reference_accel():
    """mass = 1, at rest, dt = 1: the new velocity is exactly what the reference's run() writes
    (buffer[i] = G*acc, src/nbody/nbody-run.c:22-24), and x' = x + v'."""
    n = 700
    pos = synth.plummer(n, seed=9)
    a, b = pos.copy(), np.zeros_like(pos)
    port.nbody_run(a, b, 1)                          # b = accel of state 0 (pinned to the reference)
    p, v = pos.copy(), np.zeros_like(pos)
    port.nbody_integrate(p, v, np.ones(n), 1.0, 1)
    assert np.array_equal(v, b)
    assert np.array_equal(p, pos + b)


def test_integrator_matches_a_numpy_restatement():
    n, dt, steps = 97, 0.01, 4
    pos, vel, mass = synth.plummer(n, seed=3), synth.velocities(n, seed=4), synth.masses(n, seed=5)
    p, v = pos.copy(), vel.copy()
    port.nbody_integrate(p, v, mass, dt, steps, threads=3)
    q, w = pos.copy(), vel.copy()
    for _ in range(steps):
        acc = np.zeros_like(q)
        for j in range(n):                           # sources in order: one rounding per operation
            d = q[j] - q
            s = d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2] + 0.005
            f = mass[j] * (1.0 / np.sqrt(s * s * s))
            acc += d * f[:, None]
        w = w + dt * (9.8 * acc)
        q = q + dt * w
    assert np.array_equal(p, q) and np.array_equal(v, w)


def test_integrator_energy_is_conserved_and_threads_do_not_matter():
    n = 300
    pos, vel, mass = synth.plummer(n, seed=3), synth.velocities(n, seed=4, scale=0.05), synth.masses(n, seed=5)
    k0, u0 = port.nbody_total_energy(pos, vel, mass)
    p1, v1 = pos.copy(), vel.copy()
    port.nbody_integrate(p1, v1, mass, 1e-3, 50, threads=1)
    p2, v2 = pos.copy(), vel.copy()
    port.nbody_integrate(p2, v2, mass, 1e-3, 50, threads=4)
    assert np.array_equal(p1, p2) and np.array_equal(v1, v2)
    k1, u1 = port.nbody_total_energy(p1, v1, mass)
    assert abs((k1 + u1) - (k0 + u0)) <= 1e-3 * abs(k0 + u0)
    assert k0 > 0 and u0 < 0
