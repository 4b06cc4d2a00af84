"""GPU parity tests of the pair-symmetric EXACT nbody step (csrc/nbody_sym.cuh), through the C ABI.

Every unordered pair is evaluated once and added to both bodies; each target must still receive
its sources in the order of accel() (reference src/nbody/nbody-run.c:17-25), so everything here is
BIT-identical: to the oracle at sizes it finishes in seconds, to the committed results of the
unmodified reference, and at N = 2^20 to the one-target-per-thread kernel (itself pinned to the
oracle by tests/test_nbody_gpu.py) on all 2^20 bodies.
"""
import hashlib
import os

import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth

pytestmark = [pytest.mark.gpu]
if os.environ.get("HGB_TEST_SYM", "1") == "0":      # opt-out switch for bisecting on a GPU box
    pytestmark.append(pytest.mark.skip(reason="HGB_TEST_SYM=0"))

VARIANTS = (124, 144, 146, 148, 181, 182, 183, 224, 226, 228, 241, 242, 243, 413, 421, 422, 423)   # R*100 + U*10 + CTAs per SM


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def sym_step(hgb, pos, iters, variant=1, want_prev=False):
    with hgb.nbody.NBody(pos.shape[0]) as nb:
        nb.symmetric(variant, 0).upload(pos)
        nb.step(iters)
        return nb.download(want_prev=want_prev)


@pytest.mark.parametrize("n,iters", [(1, 2), (2, 3), (3, 10), (31, 2), (63, 1), (64, 2), (65, 2), (127, 1), (128, 1),
                                     (129, 3), (200, 2), (1023, 2), (1024, 1), (1025, 2), (4097, 3), (9000, 1)])
def test_sym_ragged_sizes_vs_oracle(hgb, n, iters):
    from oracle import port
    pos = synth.uniform_cube(n, seed=n)
    a, b = pos.copy(), np.zeros_like(pos)
    port.nbody_run(a, b, iters)
    want_cur, want_prev = (a, b) if iters % 2 == 0 else (b, a)
    cur, prev = sym_step(hgb, pos, iters, want_prev=True)
    assert np.array_equal(cur, want_cur) and np.array_equal(prev, want_prev)


@pytest.mark.parametrize("variant", VARIANTS)
def test_sym_every_variant_gives_the_same_bits(hgb, variant):
    from oracle import port
    for n in (3000, 777):
        pos = synth.plummer(n, seed=7)
        want = np.zeros_like(pos)
        port.nbody_step_range(pos, want, 0, n)
        assert np.array_equal(sym_step(hgb, pos, 1, variant), want), (variant, n)


@pytest.mark.parametrize("name,gen,n,iters", [("uniform", synth.uniform_cube, 2049, 10),
                                              ("plummer", synth.plummer, 4096, 10)])
def test_sym_reference_golden_10_steps(hgb, golden, name, gen, n, iters):
    """positions after 10 steps are BIT-identical to the unmodified reference (1e-10 asked)"""
    g = golden["nbody"][f"{name}_n{n}_it{iters}"]
    cur, prev = sym_step(hgb, gen(n), iters, want_prev=True)
    assert sha(cur) == g["planets_sha256"] and sha(prev) == g["buffer_sha256"]


def test_sym_fill_planet_shipped_config(hgb, golden):
    g = golden["nbody"]["fill_planet_n3_it10"]
    cur = sym_step(hgb, synth.fill_planets(3), 10)
    assert np.array_equal(cur, np.array([float.fromhex(x) for x in g["planets"]]).reshape(3, 3))


def test_sym_out_of_range_operands_fall_back_to_the_ieee_kernel(hgb):
    """overflowing |d|^2 (t = inf) or a NaN body: the wavefront kernel flags the step and the
    ordinary kernel redoes it on the device -- results as the oracle's, NaNs included"""
    from oracle import port
    n = 2500
    pos = synth.uniform_cube(n, seed=3)
    pos[17] = [1e120, -3e150, 2e103]
    pos[1999] = [4e-200, 1e-310, -2e-308]
    for poison in (False, True):
        if poison:
            pos[1234, 1] = np.nan
        want = np.zeros_like(pos)
        port.nbody_step_range(pos, want, 0, n)
        got = sym_step(hgb, pos, 1)
        assert np.array_equal(np.isnan(got), np.isnan(want))
        ok = ~np.isnan(want)
        assert np.array_equal(got[ok], want[ok])


def test_sym_coincident_bodies_give_exact_zero(hgb):
    n = 5000
    pos = np.tile(np.array([[0.25, -3.0, 7.5]]), (n, 1))
    got = sym_step(hgb, pos, 1)
    assert np.all(got == 0.0) and not np.signbit(got).any()


def test_sym_duplicate_and_axis_aligned_bodies(hgb):
    """pairs with d = +0 in some component, duplicated bodies, huge dynamic range of factors"""
    from oracle import port
    n = 1500
    pos = synth.plummer(n, seed=11)
    pos[100:200] = pos[300:400]              # exact duplicates
    pos[500:600, 0] = pos[0, 0]              # shared x coordinate
    pos[700:720] *= 1e-9                     # a tight clump: f ~ EPS^-1.5
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    assert np.array_equal(sym_step(hgb, pos, 1), want)


def test_sym_64k_plummer_all_bodies_equal_the_standard_kernel(hgb):
    from oracle import port
    n = 65536
    pos = synth.plummer(n, seed=1234)
    targets = (synth.splitmix64(99, 512) % np.uint64(n)).astype(np.int32)
    want = port.nbody_accel_subset(pos, targets)
    with hgb.nbody.NBody(n) as nb:
        nb.symmetric(0, 0).upload(pos)
        nb.step(2)
        std = nb.download()
    got = sym_step(hgb, pos, 2)
    one = sym_step(hgb, pos, 1)
    assert np.array_equal(one[targets], want)
    assert np.array_equal(got, std)


def test_sym_full_size_1m_all_bodies_equal_the_standard_kernel(hgb):
    """N = 2^20, the benchmark size: all 2^20 results bit-equal to the standard kernel, and 1024
    random targets recomputed by the oracle"""
    from oracle import port
    n = 1 << 20
    pos = synth.uniform_cube(n, seed=1)
    targets = (synth.splitmix64(5, 1024) % np.uint64(n)).astype(np.int32)
    want = port.nbody_accel_subset(pos, targets)
    with hgb.nbody.NBody(n) as nb:
        nb.symmetric(0, 0).upload(pos)
        nb.step(1)
        std = nb.download()
    got = sym_step(hgb, pos, 1)
    assert np.array_equal(got[targets], want)
    assert np.array_equal(got, std)


def test_sym_is_ignored_where_it_does_not_apply(hgb):
    """FAST mode, the integrator extension and sharded handles keep their own kernels"""
    from oracle import port
    n = 3000
    pos = synth.plummer(n, seed=5)
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    with hgb.nbody.NBody(n, mode=hgb.NBODY_FAST) as nb:
        nb.symmetric(1, 0).upload(pos)
        nb.step(1)
        got = nb.download()
    assert np.abs(got - want).max() / np.abs(want).max() <= 1e-12
    with hgb.nbody.NBody(n) as nb:
        nb.symmetric(1, 10 ** 6).upload(pos)     # n below min_n: standard kernel
        nb.step(1)
        assert np.array_equal(nb.download(), want)
    with pytest.raises(hgb.HgbError):
        hgb.nbody.NBody(n).symmetric(999, 0)


def test_sym_default_policy(hgb):
    """new one-GPU handles pick the kernel by size (standard < 49152 <= cooperative rows < 229376 <= one warp
    per row < 655360 <= segmented first generation); an explicit hgb_nbody_tune variant of the standard kernel or FAST
    mode switch the pair-symmetric kernels off; hgb_nbody_set_symmetric(h, 1, min_n) asks for the first generation"""
    L = hgb.lib()
    if os.environ.get("HGB_NBODY_SYM") or os.environ.get("HGB_NBODY_SYM_MIN_N"):
        pytest.skip("defaults overridden by the environment")
    for n, want in ((49151, 0), (49152, 3002), (229375, 3002), (229376, 2002), (655359, 2002), (655360, 6242)):
        with hgb.nbody.NBody(n) as nb:
            assert L.hgb_nbody_symmetric(nb._h) == want, n
    with hgb.nbody.NBody(393216) as nb:
        nb.tune(128, 2, 24)
        assert L.hgb_nbody_symmetric(nb._h) == 0
    with hgb.nbody.NBody(393215) as nb:
        nb.symmetric(1, 0)
        assert L.hgb_nbody_symmetric(nb._h) == 242
    with hgb.nbody.NBody(1 << 19, mode=hgb.NBODY_FAST) as nb:
        assert L.hgb_nbody_symmetric(nb._h) == 0


def test_sym_many_steps_reuse_the_tagged_buffer(hgb):
    """epoch tags: 40 consecutive launches on one handle (and an upload in between) never see stale words"""
    from oracle import port
    n = 700
    pos = synth.uniform_cube(n, seed=21)
    a, b = pos.copy(), np.zeros_like(pos)
    port.nbody_run(a, b, 4)
    with hgb.nbody.NBody(n) as nb:
        nb.symmetric(1, 0).upload(pos)
        nb.step(37)
        nb.upload(pos)
        nb.step(4)
        assert np.array_equal(nb.download(), a)


def test_sym_epoch_counter_wraps_cleanly(hgb):
    """the launch epoch in the tags is 11 bits (1..2047): results stay exact across the wrap"""
    from oracle import port
    n = 100
    pos = synth.uniform_cube(n, seed=33)
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    with hgb.nbody.NBody(n) as nb:
        nb.symmetric(1, 0)
        for it in range(2100):
            nb.upload(pos)
            nb.step(1)
            if it % 300 == 0 or 2040 <= it <= 2052:
                assert np.array_equal(nb.download(), want), it
        assert np.array_equal(nb.download(), want)
