"""GPU parity tests for the nbody path, driven through the C ABI (libhgb200.so) and checked
against the oracle / the committed results of the unmodified reference.
EXACT mode: bit-identical.  FAST mode: 1e-12 max-norm relative for one step (stated below)."""
import ctypes as C
import hashlib

import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth

pytestmark = pytest.mark.gpu
F = float.fromhex


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def unhex(xs, shape):
    return np.array([F(x) for x in xs], dtype=np.float64).reshape(shape)


def ngpus(hgb):
    return hgb.lib().hgb_device_count()


@pytest.mark.parametrize("n,iters", [(3, 0), (3, 1), (3, 2), (3, 3), (3, 10), (5, 10), (16, 7)])
def test_run_fill_planet_golden(hgb, golden, n, iters):
    """drop-in run() on the shipped initial condition == the reference's own output"""
    g = golden["nbody"][f"fill_planet_n{n}_it{iters}"]
    planets, buffer = synth.fill_planets(n), np.zeros((n, 3))
    hgb.nbody.run(planets, buffer, n, iters, 1)
    assert np.array_equal(planets, unhex(g["planets"], (n, 3)))
    assert np.array_equal(buffer, unhex(g["buffer"], (n, 3)))


def test_run_shipped_default_prints_like_the_reference(hgb, golden):
    planets, buffer = synth.fill_planets(3), np.zeros((3, 3))
    hgb.nbody.run(planets, buffer, 3, 10, 1)
    assert ["(%g, %g, %g)" % tuple(r) for r in planets] == golden["cli"]["nbody -s 3 -n 10"]


def test_run_zero_work_touches_nothing(hgb):
    a, b = synth.uniform_cube(8), np.full((8, 3), 7.0)
    a0 = a.copy()
    hgb.nbody.run(a, b, 8, 0, 1)
    hgb.nbody.run(a, b, 0, 5, 1)
    assert np.array_equal(a, a0) and np.all(b == 7.0)


@pytest.mark.parametrize("name,gen,n,iters", [("uniform", synth.uniform_cube, 257, 4),
                                              ("plummer", synth.plummer, 1000, 3),
                                              ("uniform", synth.uniform_cube, 2049, 10),
                                              ("plummer", synth.plummer, 4096, 10)])
def test_run_synthetic_golden_10_steps(hgb, golden, name, gen, n, iters):
    """positions after up to 10 steps are BIT-identical to the reference (tolerance 1e-10 asked)"""
    g = golden["nbody"][f"{name}_n{n}_it{iters}"]
    a, b = gen(n), np.zeros((n, 3))
    hgb.nbody.run(a, b, n, iters, 1)
    assert sha(a) == g["planets_sha256"] and sha(b) == g["buffer_sha256"]


@pytest.mark.parametrize("n,iters", [(1, 2), (2, 3), (31, 2), (1023, 2), (1024, 1), (1025, 2), (4097, 3), (9000, 1)])
def test_step_ragged_sizes_vs_oracle(hgb, n, iters):
    from oracle import port
    pos = synth.uniform_cube(n, seed=n)
    a, b = pos.copy(), np.zeros_like(pos)
    port.nbody_run(a, b, iters)
    want_cur, want_prev = (a, b) if iters % 2 == 0 else (b, a)
    with hgb.nbody.NBody(n) as nb:
        nb.upload(pos)
        nb.step(iters)
        cur, prev = nb.download(want_prev=True)
    assert np.array_equal(cur, want_cur) and np.array_equal(prev, want_prev)


def test_every_kernel_variant_gives_the_same_bits(hgb):
    from oracle import port
    n = 3000
    pos = synth.plummer(n, seed=7)
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    shapes = [(b, u, w) for b in (128, 256) for u in (1, 2, 4) for w in (24, 32, 40)]
    shapes += [(128, 2, 28)]   # 64-thread CTAs measured no faster: built only with -DHGB_EXPERIMENTAL_VARIANTS
    for block, unroll, warps in shapes:
        with hgb.nbody.NBody(n) as nb:
            nb.tune(block, unroll, warps).upload(pos)
            nb.step(1)
            assert np.array_equal(nb.download(), want), (block, unroll, warps)
    # ragged counts around the tile sizes
    for m in (1, 63, 64, 65, 255, 257, 511, 513, 2999):
        sub = np.ascontiguousarray(pos[:m])
        w2 = np.zeros_like(sub)
        port.nbody_step_range(sub, w2, 0, m)
        with hgb.nbody.NBody(m) as nb:
            nb.symmetric(0, 0).upload(sub)
            nb.step(1)
            assert np.array_equal(nb.download(), w2), m


def test_overflow_and_nonfinite_operands_take_the_ieee_slow_path(hgb):
    """|d|^2 overflows for far bodies (t = inf -> factor 0) and a NaN body poisons everything
    it touches: the kernel must redo such tiles with the full IEEE sequences."""
    from oracle import port
    n = 2500
    pos = synth.uniform_cube(n, seed=3)
    pos[17] = [1e120, -3e150, 2e103]
    pos[1999] = [4e-200, 1e-310, -2e-308]       # denormal coordinates are ordinary operands
    for poison in (False, True):
        if poison:
            pos[1234, 1] = np.nan
        want = np.zeros_like(pos)
        port.nbody_step_range(pos, want, 0, n)
        with hgb.nbody.NBody(n) as nb:
            nb.upload(pos)
            nb.step(1)
            got = nb.download()
        assert np.array_equal(np.isnan(got), np.isnan(want))
        ok = ~np.isnan(want)
        assert np.array_equal(got[ok], want[ok])
        assert np.isfinite(want[0]).all() != poison


def test_inv_sqrt_sequences_match_ieee_on_hard_cases(hgb):
    mism, first = C.c_ulonglong(1), C.c_double(0)
    hgb._abi.check(hgb.lib().hgb_selftest_inv_sqrt(2_000_000_000, 2024, C.byref(mism), C.byref(first)))
    assert mism.value == 0, f"first bad operand {first.value.hex()}"


def test_subset_at_64k_plummer_vs_oracle(hgb):
    from oracle import port
    n = 65536
    pos = synth.plummer(n, seed=1234)
    targets = (synth.splitmix64(99, 1024) % np.uint64(n)).astype(np.int32)
    want = port.nbody_accel_subset(pos, targets)
    with hgb.nbody.NBody(n) as nb:
        nb.upload(pos)
        got_sub = nb.accel_subset(targets)
        nb.step(1)
        full = nb.download()
    assert np.array_equal(got_sub, want)
    assert np.array_equal(full[targets], want)


def test_full_size_1m_uniform_subsample(hgb):
    """N = 2^20 (the benchmark size): one full step on the GPU, 2048 random targets recomputed
    by the oracle against all 2^20 sources must match bit for bit."""
    from oracle import port
    n = 1 << 20
    pos = synth.uniform_cube(n, seed=1)
    targets = (synth.splitmix64(5, 2048) % np.uint64(n)).astype(np.int32)
    want = port.nbody_accel_subset(pos, targets)
    with hgb.nbody.NBody(n) as nb:
        nb.upload(pos)
        ms = nb.step(1)
        cur, prev = nb.download(want_prev=True)
    assert np.array_equal(cur[targets], want)
    assert np.array_equal(prev, pos)
    assert ms > 0


def test_coincident_bodies_give_exact_zero(hgb):
    # every pair has d = 0: the reference adds +0 * factor -> the new state is exactly +0.0
    n = 5000
    pos = np.tile(np.array([[0.25, -3.0, 7.5]]), (n, 1))
    with hgb.nbody.NBody(n) as nb:
        nb.upload(pos)
        nb.step(1)
        got = nb.download()
    assert np.all(got == 0.0) and not np.signbit(got).any()


def test_fast_mode_single_step_tolerance(hgb):
    """FAST mode (FMA + rsqrt-Newton) is NOT bit-exact: tolerance 1e-12 max-norm relative per step."""
    from oracle import port
    n = 8192
    pos = synth.plummer(n, seed=2)
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    with hgb.nbody.NBody(n, mode=hgb.NBODY_FAST) as nb:
        nb.upload(pos)
        nb.step(1)
        got = nb.download()
    rel = np.abs(got - want).max() / np.abs(want).max()
    assert rel <= 1e-12, rel


def test_energy_extension_vs_long_double(hgb):
    """extension (no reference counterpart): potential energy within 1e-12 relative"""
    from oracle import port
    pos = synth.plummer(6000, seed=2)
    want = port.nbody_energy(pos)
    with hgb.nbody.NBody(6000) as nb:
        nb.upload(pos)
        e0 = nb.energy()
        nb.step(2)
        e2 = nb.energy()
        cur = nb.download()
    assert abs(e0 - want) <= 1e-12 * abs(want)
    assert abs(e2 - port.nbody_energy(cur)) <= 1e-12 * abs(e2)


def test_errors_are_reported_not_hidden(hgb):
    with pytest.raises(hgb.HgbError) as e:
        hgb.nbody.NBody(0)
    assert e.value.code == hgb._abi.HGB_ERR_ARG
    with hgb.nbody.NBody(10) as nb:
        with pytest.raises(hgb.HgbError) as e:
            nb.step(1)                       # step before upload
        assert e.value.code == hgb._abi.HGB_ERR_STATE


# ---- multi-GPU (runs when the box has >= 2 GPUs) ------------------------------------------
@pytest.mark.parametrize("exchange", ["nccl", "p2p"])
@pytest.mark.parametrize("n,iters", [(5000, 3), (2048, 2), (70001, 2)])
def test_target_sharding_over_gpus_is_bit_identical(hgb, monkeypatch, exchange, n, iters):
    g = min(ngpus(hgb), 8)
    if g < 2:       # one device: HGB_VIRTUAL_GPUS shards on it, results stored peer-to-peer
        if exchange == "nccl":
            pytest.skip("NCCL needs >= 2 devices (virtual shards use peer stores)")
        monkeypatch.setenv("HGB_VIRTUAL_GPUS", "4")
        g = 4
    kind = hgb.EXCHANGE_NCCL if exchange == "nccl" else hgb.EXCHANGE_P2P
    from oracle import port
    pos = synth.plummer(n, seed=n)
    with hgb.nbody.NBody(n, ngpus=1) as one:
        one.upload(pos)
        one.step(iters)
        want_cur, want_prev = one.download(want_prev=True)
    # the oracle, not only the 1-GPU result: whole run for the small sets, a target sample of the
    # last step for the large one
    if n <= 5000:
        a, b = pos.copy(), np.zeros_like(pos)
        port.nbody_run(a, b, iters)
        o_cur, o_prev = (a, b) if iters % 2 == 0 else (b, a)
        assert np.array_equal(want_cur, o_cur) and np.array_equal(want_prev, o_prev)
    sample = np.random.default_rng(n).choice(n, size=min(n, 96), replace=False).astype(np.int32)
    for gg in sorted({2, g}):
        with hgb.nbody.NBody(n, ngpus=gg, exchange=kind) as nb:
            nb.upload(pos)
            nb.step(iters)
            cur, prev = nb.download(want_prev=True)
        assert np.array_equal(cur, want_cur) and np.array_equal(prev, want_prev), (exchange, gg)
        assert np.array_equal(cur[sample], port.nbody_accel_subset(prev, sample)), (exchange, gg)


def test_a_live_handle_survives_the_creation_of_a_wider_one(hgb, monkeypatch):
    """device contexts must not move when a handle over more GPUs is created (a 1-GPU handle keeps
    pointers to its stream and events): step the old handle after the new one exists, and again
    after the new one is gone; same for the cached drop-in handles."""
    from oracle import port
    monkeypatch.setenv("HGB_VIRTUAL_GPUS", "4")
    n = 1500
    pos = synth.plummer(n, seed=5)
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    one = hgb.nbody.NBody(n, ngpus=1)
    one.upload(pos)
    one.step(1)
    assert np.array_equal(one.download(), want)
    wide = hgb.nbody.NBody(n, ngpus=4, exchange=hgb.EXCHANGE_P2P)
    wide.upload(pos)
    wide.step(1)
    assert np.array_equal(wide.download(), want)
    one.upload(pos)
    one.step(1)                                   # contexts of devices 0..3 were added behind it
    assert np.array_equal(one.download(), want)
    wide.close()
    one.upload(pos)
    one.step(1)
    assert np.array_equal(one.download(), want)
    one.close()
    for g in (1, 4, 1):                           # drop-in: cached handles of different widths in turn
        a, b = pos.copy(), np.zeros_like(pos)
        hgb.nbody.set_exchange(hgb.EXCHANGE_P2P)
        try:
            hgb.nbody.run(a, b, n, 1, g)
        finally:
            hgb.nbody.set_exchange(hgb.EXCHANGE_NCCL)
        assert np.array_equal(b, want), g


def test_drop_in_with_number_of_processes_as_gpu_count(hgb):
    g = min(ngpus(hgb), 8)
    if g < 2:
        pytest.skip("needs >= 2 GPUs")
    from oracle import port
    pos = synth.uniform_cube(3001, seed=8)
    a, b = pos.copy(), np.zeros_like(pos)
    port.nbody_run(a, b, 3)
    ga, gb = pos.copy(), np.zeros_like(pos)
    hgb.nbody.run(ga, gb, 3001, 3, g)
    assert np.array_equal(ga, a) and np.array_equal(gb, b)
