"""GPU parity tests for the convolution path through the C ABI: bit-identical to the
reference (golden results of the unmodified reference + the oracle on seeded inputs),
including the reference's own gtest case, ragged/edge shapes and the 16384^2 benchmark size."""
import hashlib

import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth

pytestmark = pytest.mark.gpu
F = float.fromhex


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def conv_run(hgb, cur, prev, w, h, p, kernel, iters, ngpus=1):
    kernel = np.ascontiguousarray(kernel, dtype=np.float64)
    kh, kw = kernel.shape
    d = hgb._abi.dptr
    hgb._abi.check(hgb.lib().hgb_conv_run(d(cur), d(prev), w, h, p, d(kernel), kw, kh, iters, ngpus))


def test_run_laplace_neutralizes_image(hgb):
    """the reference's own hot-path test (test/convolutionTests.cpp:9-64), on the GPU"""
    cv = hgb.convolution
    img = cv.init_image(3, 3, 1)
    img.image[...] = [[-1.5, 0, -1], [9, -2, 3], [2, 0, 5]]
    args = cv.Args(number_of_iterations=15, number_of_processes=1, width=3, height=3)
    padded = [cv.add_padding(img, 1)]
    buffer = [cv.init_padded_image(3, 3, 1)]
    cv.copy_padded_image(padded[0], buffer[0])
    laplace = cv.init_image(3, 3, 0)
    laplace.image[1, 1] = -1
    laplace.image[0, 1] = laplace.image[2, 1] = laplace.image[1, 0] = laplace.image[1, 2] = 0.25
    assert cv.get_checksum(img) == 14.5
    cv.run_on_padded_image(padded, laplace, args, buffer)
    result = cv.remove_padding(padded[0])
    assert cv.get_checksum(result) == 0


def test_gtest_case_matches_reference_bits(hgb, golden):
    g = golden["conv"]["gtest_laplace_15"]
    cur = synth.add_padding(np.array([[-1.5, 0, -1], [9, -2, 3], [2, 0, 5]]), 1)
    prev = cur.copy()
    conv_run(hgb, cur, prev, 3, 3, 1, synth.LAPLACE_TXT, 15)
    assert [float(v).hex() for v in cur.ravel()] == g["cur"]
    assert [float(v).hex() for v in prev.ravel()] == g["prev"]


def _golden_case(hgb, golden, key, dense, kernel):
    g = golden["conv"][key]
    p, it = g["p"], g["iters"]
    h, w = dense.shape
    cur = synth.add_padding(dense, p)
    prev = np.zeros_like(cur)
    conv_run(hgb, cur, prev, w, h, p, kernel, it)
    assert sha(cur) == g["cur_sha256"], key
    assert sha(prev) == g["prev_sha256"], key
    # device-side sequential checksum == reference get_checksum
    with hgb.convolution.Conv2D(w, h, p) as cv:
        cv.set_kernel(kernel).upload(synth.add_padding(dense, p))
        cv.step(it)
        cs = cv.checksum()
    assert cs == F(g["checksum"]) and "%f" % cs == g["checksum_printed"]


@pytest.mark.parametrize("it", [1, 10, 100])
def test_shipped_image_kernel_txt(hgb, golden, shipped_image, it):
    _golden_case(hgb, golden, f"input_kernel_txt_n{it}", shipped_image, synth.KERNEL_TXT)


@pytest.mark.parametrize("it", [1, 10, 100])
def test_shipped_image_laplace_txt(hgb, golden, shipped_image, it):
    _golden_case(hgb, golden, f"input_laplace_txt_n{it}", shipped_image, synth.LAPLACE_TXT)


def test_shipped_image_default_kernel(hgb, golden, shipped_image):
    _golden_case(hgb, golden, "input_default5x5_n10", shipped_image, synth.default_kernel())


def test_shipped_float_image(hgb, golden, shipped_image):
    _golden_case(hgb, golden, "txt_kernel_txt_n10", shipped_image / 256.0, synth.KERNEL_TXT)


@pytest.mark.parametrize("w,h,k,it", [(33, 20, 7, 4), (300, 70, 3, 3), (513, 260, 5, 2),
                                      (1024, 768, 3, 5), (640, 480, 7, 3)])
def test_synthetic_float_golden(hgb, golden, w, h, k, it):
    _golden_case(hgb, golden, f"synth_float_{w}x{h}_k{k}_n{it}", synth.image_float(w, h), synth.kernel_dense(k))


@pytest.mark.parametrize("k,it", [(3, 4), (7, 2)])
def test_synthetic_int_golden(hgb, golden, k, it):
    _golden_case(hgb, golden, f"synth_int_512x384_k{k}_n{it}", synth.image_int(512, 384), synth.kernel_dense(k))


RAGGED = [  # (w, h, p, kernel shape (kh, kw), iterations)
    (1, 1, 1, (3, 3), 3), (2, 1, 1, (3, 3), 2), (1, 9, 3, (7, 7), 2), (5, 2, 3, (7, 7), 3),
    (255, 3, 1, (3, 3), 2), (256, 4, 1, (3, 3), 2), (257, 5, 1, (3, 3), 2), (511, 130, 2, (5, 5), 2),
    (769, 129, 3, (7, 7), 2), (1000, 257, 1, (3, 3), 3),
    (40, 30, 2, (3, 5), 2), (40, 30, 2, (5, 3), 2), (40, 30, 2, (4, 4), 2), (40, 30, 3, (3, 3), 2),
    (64, 64, 0, (1, 1), 3), (90, 17, 4, (9, 9), 2), (33, 12, 2, (1, 5), 2), (33, 12, 2, (2, 1), 2),
]


@pytest.mark.parametrize("w,h,p,kshape,it", RAGGED)
def test_ragged_shapes_and_kernels_vs_oracle(hgb, w, h, p, kshape, it):
    from oracle import port
    kh, kw = kshape
    kern = (2.0 * synth.uniform01(77, kh * kw) - 1.0).reshape(kh, kw)
    base = synth.add_padding(synth.image_float(w, h, seed=w * 31 + h), p)
    want_c, want_p = base.copy(), np.zeros_like(base)
    port.conv_run(want_c, want_p, w, h, p, kern, it)
    got_c, got_p = base.copy(), np.full_like(base, -1.0)
    conv_run(hgb, got_c, got_p, w, h, p, kern, it)
    assert np.array_equal(got_c, want_c) and np.array_equal(got_p, want_p)


def test_pad_columns_are_read_as_given_on_the_first_iteration(hgb):
    """the reference reads the caller's pad columns as they are (they need not be edge copies)"""
    from oracle import port
    w, h, p = 37, 9, 2
    base = (synth.uniform01(8, h * (w + 2 * p)) * 10).reshape(h, w + 2 * p)   # arbitrary pad columns
    kern = synth.kernel_dense(5)
    want_c, want_p = base.copy(), np.zeros_like(base)
    port.conv_run(want_c, want_p, w, h, p, kern, 2)
    got_c, got_p = base.copy(), np.zeros_like(base)
    conv_run(hgb, got_c, got_p, w, h, p, kern, 2)
    assert np.array_equal(got_c, want_c) and np.array_equal(got_p, want_p)


def test_iteration_count_semantics(hgb):
    w, h, p = 20, 10, 1
    base = synth.add_padding(synth.image_float(w, h), p)
    cur, prev = base.copy(), np.full_like(base, 5.0)
    conv_run(hgb, cur, prev, w, h, p, synth.LAPLACE_TXT, 0)
    assert np.array_equal(cur, base) and np.all(prev == 5.0)        # zero iterations: untouched
    conv_run(hgb, cur, prev, w, h, p, synth.LAPLACE_TXT, 1)
    assert np.array_equal(prev, base) and not np.array_equal(cur, base)  # one: buffer shows state 0


def test_both_kernel_variants_agree(hgb):
    w, h = 700, 300
    for k in (3, 5, 7):
        p = k // 2
        img = synth.add_padding(synth.image_float(w, h, seed=k), p)
        outs = []
        for variant, band in [(0, 0), (0, 7), (0, 64), (1, 0)]:
            with hgb.convolution.Conv2D(w, h, p) as cv:
                cv.set_kernel(synth.kernel_dense(k)).tune(variant, band).upload(img)
                cv.step(3)
                outs.append(cv.download())
        for o in outs[1:]:
            assert np.array_equal(o, outs[0])


def test_signed_zero_and_denormal_products(hgb):
    """val starts at +0.0: an all-(-0.0) sum must come out as +0.0; denormal products are exact ops"""
    from oracle import port
    w, h, p = 64, 8, 1
    img = np.zeros((h, w))
    img[2, :] = -0.0
    img[3, ::2] = 5e-324
    img[4, :] = -1e-310
    base = synth.add_padding(img, p)
    kern = np.array([[0.0, -0.0, 0.0], [1e-10, 1.0, -1e-10], [0.0, 0.5, 0.0]])
    want_c, want_p = base.copy(), np.zeros_like(base)
    port.conv_run(want_c, want_p, w, h, p, kern, 2)
    got_c, got_p = base.copy(), np.zeros_like(base)
    conv_run(hgb, got_c, got_p, w, h, p, kern, 2)
    assert np.array_equal(got_c, want_c) and np.array_equal(np.signbit(got_c), np.signbit(want_c))


def test_argument_errors(hgb):
    cur, prev = np.zeros((4, 6)), np.zeros((4, 6))
    E = hgb._abi
    with pytest.raises(hgb.HgbError) as e:     # 5x5 kernel needs padding 2: the reference would index out of its rows
        conv_run(hgb, cur, prev, 4, 4, 1, np.ones((5, 5)), 1)
    assert e.value.code == E.HGB_ERR_ARG
    with pytest.raises(hgb.HgbError):
        conv_run(hgb, cur, prev, 0, 4, 1, np.ones((3, 3)), 1)
    with pytest.raises(hgb.HgbError):
        conv_run(hgb, cur, prev, 4, 4, 1, np.ones((3, 3)), -1)
    with hgb.convolution.Conv2D(4, 4, 1) as cv:
        with pytest.raises(hgb.HgbError) as e:
            cv.step(1)
        assert e.value.code == E.HGB_ERR_STATE


@pytest.mark.parametrize("k", [3, 7])
def test_full_size_16k_bands_vs_oracle(hgb, k):
    """16384 x 16384 (the benchmark size): one iteration on the GPU; three 256-row bands plus the
    top and bottom edges recomputed by the oracle must match bit for bit, and an exact-arithmetic
    property must hold on the whole image: integer pixels x power-of-two taps summing to 1 keep a
    constant image constant."""
    from oracle import port
    n = 16384
    p = k // 2
    kern = synth.kernel_dense(k)
    dense = synth.image_int(n, n, seed=4)
    inp = synth.add_padding(dense, p)
    del dense
    with hgb.convolution.Conv2D(n, n, p) as cv:
        cv.set_kernel(kern).upload(inp)
        cv.step(1)
        got = cv.download()
    for r0 in (0, 5000, 9999, n - 256):
        want = _oracle_rows(port, inp, n, n, p, kern, r0, r0 + 256)
        assert np.array_equal(got[r0:r0 + 256], want), (k, r0)


@pytest.mark.parametrize("it", [3, 6])
def test_full_size_16k_fused_passes_vs_oracle(hgb, it):
    """16384 x 16384 float image, 3x3, `it` iterations (two per pass + the plain final ones): bands
    recomputed by the oracle with an `it`-row margin (pollution from the sub-image's artificial
    edges travels one row per iteration), incl. the true top and bottom edges; and the whole image
    against the one-iteration-per-pass kernel."""
    from oracle import port
    n, p = 16384, 1
    kern = synth.kernel_dense(3)
    dense = synth.image_float(n, n, seed=3)
    with hgb.convolution.Conv2D(n, n, p) as cv:
        cv.set_kernel(kern).upload_dense(dense)
        cv.step(it)
        got = cv.download()
        cv.tune(4, 0).upload_dense(dense)                      # bit 2: plain passes only
        cv.step(it)
        plain = cv.download()
    assert np.array_equal(got, plain)
    del plain
    for r0 in (0, 7777, n - 256):
        lo, hi = max(0, r0 - it), min(n, r0 + 256 + it)
        a = synth.add_padding(dense[lo:hi], p)
        b = np.zeros_like(a)
        port.conv_run(a, b, n, hi - lo, p, kern, it)
        assert np.array_equal(got[r0:r0 + 256], a[r0 - lo:r0 - lo + 256]), r0


def _oracle_rows(port, inp, w, h, p, kern, r0, r1):
    """oracle output rows [r0, r1) without allocating a second full image: run it on the
    sub-image [r0-p, r1+p) whose first/last rows are only touched through clamping when they
    are the true image edges."""
    lo, hi = max(0, r0 - p), min(h, r1 + p)
    sub = np.ascontiguousarray(inp[lo:hi])
    out = np.zeros_like(sub)
    port.conv_step_rows(sub, out, w, hi - lo, p, kern, r0 - lo, r1 - lo)
    return out[r0 - lo:r1 - lo]


def test_constant_image_stays_constant_at_full_size(hgb):
    n, p = 16384, 2
    base = np.full((n, n + 2 * p), 3.0)
    with hgb.convolution.Conv2D(n, n, p) as cv:
        cv.set_kernel(synth.KERNEL_TXT).upload(base)     # taps are powers of two summing to 1
        cv.step(4)
        got = cv.download()
    assert np.all(got == 3.0)


# ---- multi-GPU: real devices when the box has them, else HGB_VIRTUAL_GPUS shards on one device ----
def _gpu_counts(hgb, monkeypatch, exchange):
    g = min(hgb.lib().hgb_device_count(), 8)
    if g < 2:
        if exchange == "nccl":
            pytest.skip("NCCL needs >= 2 devices (virtual shards use peer copies)")
        monkeypatch.setenv("HGB_VIRTUAL_GPUS", "4")
        return [2, 3, 4]
    return sorted({2, g})


@pytest.mark.parametrize("exchange", ["nccl", "p2p"])
@pytest.mark.parametrize("depth", [0, 1, 3])
@pytest.mark.parametrize("w,h,k,it", [(300, 70, 3, 4), (513, 260, 5, 3), (640, 33, 7, 3), (64, 5, 7, 2),
                                      (300, 1100, 3, 21), (513, 1300, 7, 9), (260, 2049, 5, 10)])  # tall: overlapped path
def test_row_bands_over_gpus_are_bit_identical(hgb, monkeypatch, exchange, depth, w, h, k, it):
    counts = _gpu_counts(hgb, monkeypatch, exchange)
    p = k // 2
    kind = hgb.EXCHANGE_NCCL if exchange == "nccl" else hgb.EXCHANGE_P2P
    from oracle import port
    img = synth.add_padding(synth.image_float(w, h, seed=w), p)
    with hgb.convolution.Conv2D(w, h, p, ngpus=1) as one:
        one.set_kernel(synth.kernel_dense(k)).upload(img)
        one.step(it)
        want_c, want_p = one.download(want_prev=True)
    o_c, o_p = img.copy(), np.zeros_like(img)                 # the oracle, not only the 1-GPU result
    port.conv_run(o_c, o_p, w, h, p, synth.kernel_dense(k), it)
    assert np.array_equal(want_c, o_c) and np.array_equal(want_p, o_p)
    for gg in counts:
        with hgb.convolution.Conv2D(w, h, p, ngpus=gg) as cv:
            cv.set_exchange(kind).set_kernel(synth.kernel_dense(k)).upload(img)
            d = cv.set_depth(depth)
            assert d >= 1 and (depth == 0 or d <= depth)
            cv.step(it)
            cur, prev = cv.download(want_prev=True)
        assert np.array_equal(cur, want_c) and np.array_equal(prev, want_p), (exchange, gg, d)


def test_depth_can_change_between_steps_and_split_steps_agree(hgb, monkeypatch):
    counts = _gpu_counts(hgb, monkeypatch, "p2p")
    w, h, k, p = 384, 1500, 3, 1
    img = synth.add_padding(synth.image_float(w, h, seed=9), p)
    with hgb.convolution.Conv2D(w, h, p, ngpus=1) as one:
        one.set_kernel(synth.kernel_dense(k)).upload(img)
        one.step(17)
        want = one.download()
    with hgb.convolution.Conv2D(w, h, p, ngpus=counts[-1]) as cv:
        cv.set_exchange(hgb.EXCHANGE_P2P).set_kernel(synth.kernel_dense(k)).upload(img)
        cv.set_depth(2)
        cv.step(5)
        cv.set_depth(16)          # deeper than the halos exchanged so far: step re-exchanges first
        cv.step(11)
        cv.set_depth(1)
        cv.step(1)
        assert np.array_equal(cv.download(), want)


# ---- 3x3 two-iterations-per-pass kernel (temporal blocking): bit-identical to separate passes ----
FUSED = [(1, 1), (2, 1), (1, 2), (3, 3), (2, 7), (5, 2), (253, 4), (254, 3), (255, 5), (256, 6), (257, 9),
         (508, 2), (509, 33), (511, 70), (1000, 37), (1271, 130)]


@pytest.mark.parametrize("w,h", FUSED)
@pytest.mark.parametrize("it,band,variant", [(3, 0, 0), (4, 0, 8), (6, 1, 0), (7, 2, 8), (5, 3, 0), (8, 5, 0), (5, 4, 8),
                                             (3, 0, 16), (6, 1, 16), (7, 2, 16), (8, 5, 16)])
def test_fused_3x3_matches_oracle_on_edges_and_strip_seams(hgb, w, h, it, band, variant):
    """odd/even widths around the 254-column strip seam, tiny heights (top and bottom clamp in
    the same chunk), arbitrary pad columns on input, odd/even iteration counts, chunk heights"""
    from oracle import port
    base = (synth.uniform01(w * 131 + h, h * (w + 2)) * 4 - 2).reshape(h, w + 2).copy()   # pads as given
    kern = synth.kernel_dense(3, seed=w + it)
    want_c, want_p = base.copy(), np.zeros_like(base)
    port.conv_run(want_c, want_p, w, h, 1, kern, it)
    with hgb.convolution.Conv2D(w, h, 1) as cv:
        cv.set_kernel(kern).tune(variant, band).upload(base)     # variant 0: 2 columns per thread, 8: 4, 16: warp-independent
        cv.step(it)
        got_c, got_p = cv.download(want_prev=True)
    assert np.array_equal(got_c, want_c)
    assert np.array_equal(got_p, want_p)          # state it-1 exists: the last iteration is a plain pass


def test_fused_3x3_equals_unfused_at_size(hgb):
    w, h, it = 4100, 2100, 9
    img = synth.image_float(w, h, seed=5)
    outs = []
    for variant in (0, 8, 16, 4):                  # 2 / 4 columns per thread, warp-independent, bit 2: no fusion
        with hgb.convolution.Conv2D(w, h, 1) as cv:
            cv.set_kernel(synth.kernel_dense(3)).tune(variant, 0).upload_dense(img)
            cv.step(it)
            outs.append(cv.download(want_prev=True))
    for o in outs[1:]:
        assert np.array_equal(outs[0][0], o[0]) and np.array_equal(outs[0][1], o[1])


def test_fused_3x3_split_steps_and_special_values(hgb):
    """step(a) then step(b) == step(a+b); signed zeros / denormals / huge values pass through both stages"""
    w, h = 300, 40
    img = synth.image_float(w, h, seed=9) - 0.5
    img[0, 0], img[3, 7], img[h - 1, w - 1], img[5, 5], img[9, 250:260] = -0.0, 5e-324, 1e300, -1e-310, 0.0
    kern = synth.kernel_dense(3)
    with hgb.convolution.Conv2D(w, h, 1) as cv:
        cv.set_kernel(kern).upload_dense(img)
        cv.step(11)
        want = cv.download(want_prev=True)
        cv.upload_dense(img)
        cv.step(4)
        cv.step(1)
        cv.step(6)
        got = cv.download(want_prev=True)
        cv.tune(4, 0).upload_dense(img)
        cv.step(11)
        plain = cv.download(want_prev=True)
    for a, b, c in zip(want, got, plain):
        assert a.tobytes() == b.tobytes() == c.tobytes()


# ---- FAST mode (fused multiply-add taps): opt-in, floating-point images only ----------------------
@pytest.mark.parametrize("w,h,kshape,it", [(300, 70, (3, 3), 10), (513, 260, (5, 5), 10), (640, 133, (7, 7), 10),
                                           (257, 64, (3, 5), 6), (1024, 1024, (3, 3), 10), (1024, 512, (7, 7), 10)])
def test_fast_mode_is_within_1e_12_of_the_reference_on_float_images(hgb, w, h, kshape, it):
    """north_star tolerance for floating-point images: 1e-12 relative (max-norm) after 10 iterations;
    the EXACT mode stays the parity path, FAST only contracts each tap's multiply-add into one FMA
    (reference arithmetic relaxed: src/convolution/convolution-run.c:71-77)"""
    from oracle import port
    kh, kw = kshape
    p = max(kh, kw) // 2
    kern = synth.kernel_dense(max(kh, kw))[:kh, :kw].copy()
    kern /= np.abs(kern).sum()                           # keep 10 iterations in range
    img = synth.add_padding(synth.image_float(w, h, seed=w + h), p)
    want_c, want_p = img.copy(), np.zeros_like(img)
    port.conv_run(want_c, want_p, w, h, p, kern, it)
    with hgb.convolution.Conv2D(w, h, p) as cv:
        cv.set_mode(hgb.CONV_FAST).set_kernel(kern).upload(img)
        assert cv.mode_in_effect() == hgb.CONV_FAST
        cv.step(it)
        cur, prev = cv.download(want_prev=True)
    tol = 1e-12                                          # the stated tolerance
    assert np.abs(cur - want_c).max() <= tol * np.abs(want_c).max()
    assert np.abs(prev - want_p).max() <= tol * np.abs(want_p).max()
    assert not np.array_equal(cur, want_c) or it == 0    # it really is the other arithmetic


@pytest.mark.parametrize("k", [3, 5, 7])
def test_fast_mode_keeps_integer_images_bit_exact(hgb, k):
    """an integer-valued image is classified at upload and keeps the exact kernels"""
    from oracle import port
    w, h, p, it = 400, 90, k // 2, 5
    kern = synth.kernel_dense(k)
    img = synth.add_padding(synth.image_int(w, h, seed=4), p)
    want_c, want_p = img.copy(), np.zeros_like(img)
    port.conv_run(want_c, want_p, w, h, p, kern, it)
    with hgb.convolution.Conv2D(w, h, p) as cv:
        cv.set_mode(hgb.CONV_FAST).set_kernel(kern).upload(img)
        assert cv.mode_in_effect() == hgb.CONV_EXACT
        cv.step(it)
        cur, prev = cv.download(want_prev=True)
        assert np.array_equal(cur, want_c) and np.array_equal(prev, want_p)
        img[17, 33] += 0.5                               # one non-integer pixel: a floating-point image
        cv.upload(img)
        assert cv.mode_in_effect() == hgb.CONV_FAST
        with pytest.raises(hgb.HgbError):                # the mode is fixed at upload
            hgb.convolution.Conv2D(8, 8, 1).set_kernel(synth.kernel_dense(3)).upload(np.zeros((8, 10))).set_mode(hgb.CONV_FAST)


def test_fast_mode_row_bands_agree_with_one_gpu(hgb, monkeypatch):
    """same arithmetic on every band: FAST over row bands is bit-identical to FAST on one GPU"""
    counts = _gpu_counts(hgb, monkeypatch, "p2p")
    w, h, k, p, it = 513, 1300, 3, 1, 9
    kern = synth.kernel_dense(k) / 5.0
    img = synth.add_padding(synth.image_float(w, h, seed=77), p)
    with hgb.convolution.Conv2D(w, h, p, ngpus=1) as one:
        one.set_mode(hgb.CONV_FAST).set_kernel(kern).upload(img)
        one.step(it)
        want = one.download()
    for g in counts:
        with hgb.convolution.Conv2D(w, h, p, ngpus=g) as cv:
            cv.set_exchange(hgb.EXCHANGE_P2P).set_mode(hgb.CONV_FAST).set_kernel(kern).upload(img)
            cv.step(it)
            assert np.array_equal(cv.download(), want), g


# ---- hgb_conv_run streamed in row chunks (PCIe copies under the iterations) --------------------------------
@pytest.mark.parametrize("w,h,k,it,chunk", [(300, 1100, 3, 4, 128), (513, 1300, 7, 3, 200), (260, 2049, 5, 5, 300),
                                            (64, 700, 3, 11, 100), (1000, 900, 3, 2, 250)])
def test_streamed_run_is_bit_identical_to_the_oracle(hgb, w, h, k, it, chunk):
    from oracle import port
    p = k // 2
    kern = synth.kernel_dense(k)
    img = synth.add_padding(synth.image_float(w, h, seed=w + h + k), p)
    want_c, want_p = img.copy(), np.zeros_like(img)
    port.conv_run(want_c, want_p, w, h, p, kern, it)
    try:
        hgb.convolution.set_streaming(chunk)
        cur, prev = img.copy(), np.full_like(img, -1.0)       # pageable memory: registered by the library
        hgb.convolution.run_flat(cur, prev, w, h, p, kern, it)
        assert np.array_equal(cur, want_c) and np.array_equal(prev, want_p)
        cur2 = img.copy()
        hgb.convolution.run_flat(cur2, None, w, h, p, kern, it)    # state n-1 not wanted
        assert np.array_equal(cur2, want_c)
        hgb.convolution.set_streaming(0)                           # one piece: same bits
        cur3, prev3 = img.copy(), np.zeros_like(img)
        hgb.convolution.run_flat(cur3, prev3, w, h, p, kern, it)
        assert np.array_equal(cur3, want_c) and np.array_equal(prev3, want_p)
    finally:
        hgb.convolution.set_streaming(1)
        hgb.convolution.unpin()


def test_streamed_run_keeps_the_drop_in_semantics(hgb):
    """run_on_padded_image through the streamed path: newest state in padded_img[0], the one before in buffer[0]"""
    from oracle import port
    w, h, p, it = 200, 1500, 1, 3
    kern = synth.kernel_dense(3)
    dense = synth.image_float(w, h, seed=77)
    img = synth.add_padding(dense, p)
    want_c, want_p = img.copy(), np.zeros_like(img)
    port.conv_run(want_c, want_p, w, h, p, kern, it)
    try:
        hgb.convolution.set_streaming(100)
        cur, prev = img.copy(), np.zeros_like(img)
        hgb.convolution.run_flat(cur, prev, w, h, p, kern, it)
        assert np.array_equal(cur, want_c) and np.array_equal(prev, want_p)
    finally:
        hgb.convolution.set_streaming(1)
        hgb.convolution.unpin()


def test_streamed_runs_on_buffers_freed_between_calls(hgb):
    """the reference's run_on_padded_image callers (and the shim) free cur/prev as soon as the call is back and
    allocate fresh ones for the next image; malloc tends to hand the same address out again.  No unpin between
    the calls: a page-lock registration that outlived the first buffers would make the second call copy through
    dead mappings."""
    from oracle import port
    w, h, p, it = 700, 3000, 1, 2                                  # 2 x 16.8 MB: mmap'ed, unmapped on free
    kern = synth.kernel_dense(3)
    seen = set()
    try:
        hgb.convolution.set_streaming(200)
        for seed in (5, 6, 7):
            img = synth.add_padding(synth.image_float(w, h, seed=seed), p)
            want_c, want_p = img.copy(), np.zeros_like(img)
            port.conv_run(want_c, want_p, w, h, p, kern, it)
            cur, prev = img.copy(), np.full_like(img, -1.0)
            seen.add((cur.ctypes.data, prev.ctypes.data))
            hgb.convolution.run_flat(cur, prev, w, h, p, kern, it)
            assert np.array_equal(cur, want_c) and np.array_equal(prev, want_p), seed
            del cur, prev, img, want_c, want_p
    finally:
        hgb.convolution.set_streaming(1)
