"""GPU parity tests of the second-generation pair-symmetric EXACT nbody step (csrc/nbody_tri.cuh),
through the C ABI.  Every unordered pair is evaluated once; each target still receives its sources in
the order of accel() (reference src/nbody/nbody-run.c:17-25), so everything here is BIT-identical: to
the oracle at sizes it finishes in seconds, to the committed results of the unmodified reference, and at
N = 2^20 to the one-target-per-thread kernel on all 2^20 bodies -- on one GPU and with the rows dealt
over several (real devices when the box has them, else HGB_VIRTUAL_GPUS shards on one device).
"""
import hashlib

import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth

pytestmark = [pytest.mark.gpu]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def tri_step(hgb, pos, iters, variant=2, seg=0, ngpus=1, want_prev=False, rows=0, weight=0):
    kind = hgb.EXCHANGE_P2P if ngpus > 1 else hgb.EXCHANGE_NCCL
    with hgb.nbody.NBody(pos.shape[0], ngpus=ngpus, exchange=kind) as nb:
        nb.symmetric(variant, 0).segment(seg, rows, weight).upload(pos)
        assert hgb.lib().hgb_nbody_symmetric(nb._h) >= 2001
        nb.step(iters)
        return nb.download(want_prev=want_prev)


def oracle_run(pos, iters):
    from oracle import port
    a, b = pos.copy(), np.zeros_like(pos)
    port.nbody_run(a, b, iters)
    return (a, b) if iters % 2 == 0 else (b, a)


@pytest.mark.parametrize("n,iters,seg", [(1, 2, 32), (2, 3, 32), (3, 10, 0), (31, 2, 32), (32, 1, 32), (33, 2, 32),
                                         (63, 1, 64), (64, 2, 32), (65, 2, 64), (127, 1, 32), (129, 3, 96),
                                         (200, 2, 64), (1023, 2, 32), (1024, 1, 256), (1025, 2, 0),
                                         (4097, 3, 512), (9000, 1, 1024), (9000, 2, 0)])
def test_tri_ragged_sizes_vs_oracle(hgb, n, iters, seg):
    pos = synth.uniform_cube(n, seed=n)
    want_cur, want_prev = oracle_run(pos, iters)
    cur, prev = tri_step(hgb, pos, iters, seg=seg, want_prev=True)
    assert np.array_equal(cur, want_cur) and np.array_equal(prev, want_prev)


@pytest.mark.parametrize("variant", [2001, 2002, 2003, 2011, 2012, 2013])   # 202x (fused row pass, +0.3 %): experimental builds only
@pytest.mark.parametrize("seg,rows,weight", [(64, 1, 1), (64, 3, 2), (2048, 16, 16), (96, 256, 1), (0, 0, 0)])
def test_tri_every_variant_and_schedule_gives_the_same_bits(hgb, variant, seg, rows, weight):
    from oracle import port
    for n in (3000, 777):
        pos = synth.plummer(n, seed=7)
        want = np.zeros_like(pos)
        port.nbody_step_range(pos, want, 0, n)
        got = tri_step(hgb, pos, 1, variant, seg, rows=rows, weight=weight)
        assert np.array_equal(got, want), (variant, seg, rows, weight, n)


@pytest.mark.parametrize("name,gen,n,iters", [("uniform", synth.uniform_cube, 2049, 10),
                                              ("plummer", synth.plummer, 4096, 10)])
def test_tri_reference_golden_10_steps(hgb, golden, name, gen, n, iters):
    """positions after 10 steps are BIT-identical to the unmodified reference (1e-10 asked)"""
    g = golden["nbody"][f"{name}_n{n}_it{iters}"]
    cur, prev = tri_step(hgb, gen(n), iters, seg=256, want_prev=True)
    assert sha(cur) == g["planets_sha256"] and sha(prev) == g["buffer_sha256"]


def test_tri_out_of_range_operands_fall_back_to_the_ieee_kernel(hgb):
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
        got = tri_step(hgb, pos, 1, seg=128)
        assert np.array_equal(np.isnan(got), np.isnan(want))
        ok = ~np.isnan(want)
        assert np.array_equal(got[ok], want[ok])


def test_tri_duplicate_coincident_and_clumped_bodies(hgb):
    from oracle import port
    n = 1500
    pos = synth.plummer(n, seed=11)
    pos[100:200] = pos[300:400]              # exact duplicates
    pos[500:600, 0] = pos[0, 0]              # shared x coordinate
    pos[700:720] *= 1e-9                     # a tight clump: f ~ EPS^-1.5
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    assert np.array_equal(tri_step(hgb, pos, 1, seg=96), want)
    same = np.tile(np.array([[0.25, -3.0, 7.5]]), (5000, 1))
    got = tri_step(hgb, same, 1)
    assert np.all(got == 0.0) and not np.signbit(got).any()


def test_tri_many_steps_and_epoch_wrap(hgb):
    """double-buffered tagged words, 11-bit launch epoch: 40 launches on one handle, an upload in between,
    and the wrap of the epoch counter never see stale words"""
    n = 700
    pos = synth.uniform_cube(n, seed=21)
    want4, _ = oracle_run(pos, 4)
    want1, _ = oracle_run(pos, 1)
    with hgb.nbody.NBody(n) as nb:
        nb.symmetric(2, 0).segment(64, 2, 3).upload(pos)
        nb.step(37)
        nb.upload(pos)
        nb.step(4)
        assert np.array_equal(nb.download(), want4)
        for it in range(2100):
            nb.upload(pos)
            nb.step(1)
            if it % 300 == 0 or 2000 <= it <= 2012:
                assert np.array_equal(nb.download(), want1), it


def _gpu_counts(hgb, monkeypatch):
    g = min(hgb.lib().hgb_device_count(), 8)
    if g < 2:
        monkeypatch.setenv("HGB_VIRTUAL_GPUS", "4")
        return [2, 3, 4]
    return sorted({2, g})


@pytest.mark.parametrize("n,iters,seg", [(5000, 3, 64), (2048, 2, 32), (70001, 2, 1024), (100, 2, 32), (40000, 1, 0)])
def test_tri_rows_dealt_over_gpus_are_bit_identical(hgb, monkeypatch, n, iters, seg):
    from oracle import port
    counts = _gpu_counts(hgb, monkeypatch)
    pos = synth.plummer(n, seed=n)
    one_cur, one_prev = tri_step(hgb, pos, iters, seg=seg, want_prev=True)
    if n <= 5000:
        want_cur, want_prev = oracle_run(pos, iters)
        assert np.array_equal(one_cur, want_cur) and np.array_equal(one_prev, want_prev)
    sample = np.random.default_rng(n).choice(n, size=min(n, 96), replace=False).astype(np.int32)
    assert np.array_equal(one_cur[sample], port.nbody_accel_subset(one_prev, sample))
    for g in counts:
        for rows in (0, 1, 5):
            cur, prev = tri_step(hgb, pos, iters, seg=seg, ngpus=g, want_prev=True, rows=rows, weight=rows)
            assert np.array_equal(cur, one_cur) and np.array_equal(prev, one_prev), (g, rows)


def test_tri_64k_plummer_all_bodies_equal_the_standard_kernel(hgb):
    from oracle import port
    n = 65536
    pos = synth.plummer(n, seed=1234)
    targets = (synth.splitmix64(99, 512) % np.uint64(n)).astype(np.int32)
    want = port.nbody_accel_subset(pos, targets)
    with hgb.nbody.NBody(n) as nb:
        nb.symmetric(0, 0).upload(pos)
        nb.step(2)
        std = nb.download()
    assert np.array_equal(tri_step(hgb, pos, 1)[targets], want)
    assert np.array_equal(tri_step(hgb, pos, 2), std)


def test_tri_full_size_1m_all_bodies_equal_the_standard_kernel(hgb):
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
    got = tri_step(hgb, pos, 1)
    assert np.array_equal(got[targets], want)
    assert np.array_equal(got, std)
