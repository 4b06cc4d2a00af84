"""GPU parity tests of the segmented first-generation pair-symmetric step (csrc/nbody_seg.cuh: the systolic warp
of nbody_sym_kernel on the rows-by-segments schedule), through the C ABI.  Every target still receives its
sources in the order of accel() (reference src/nbody/nbody-run.c:17-25): everything is BIT-identical -- to the
oracle at sizes it finishes in seconds, to the committed results of the unmodified reference, to the standard
kernel on all bodies at 64k and 2^20."""
import hashlib

import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth

pytestmark = [pytest.mark.gpu]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def seg_step(hgb, pos, iters, variant=4242, seg=0, rows=0, weight=0, want_prev=False, ngpus=1, deal=0):
    kind = hgb.EXCHANGE_P2P if ngpus > 1 else hgb.EXCHANGE_NCCL
    with hgb.nbody.NBody(pos.shape[0], ngpus=ngpus, exchange=kind) as nb:
        nb.symmetric(variant, 0).segment(seg, rows, weight).deal(deal).upload(pos)
        assert hgb.lib().hgb_nbody_symmetric(nb._h) == variant
        assert nb.describe()[0].startswith("nbody_seg_kernel")
        nb.step(iters)
        return nb.download(want_prev=want_prev)


def oracle_run(pos, iters):
    from oracle import port
    a, b = pos.copy(), np.zeros_like(pos)
    port.nbody_run(a, b, iters)
    return (a, b) if iters % 2 == 0 else (b, a)


@pytest.mark.parametrize("n,iters,seg", [(1, 2, 32), (3, 10, 0), (31, 2, 32), (32, 1, 32), (33, 2, 32), (63, 2, 32), (64, 2, 64),
                                         (65, 2, 64), (127, 1, 32), (129, 3, 96), (200, 2, 64), (1023, 2, 32), (1025, 2, 0),
                                         (2000, 1, 160), (4097, 3, 512), (9000, 1, 1024), (9000, 2, 0)])
@pytest.mark.parametrize("variant", [4242, 6242])
def test_seg_ragged_sizes_vs_oracle(hgb, variant, n, iters, seg):
    pos = synth.uniform_cube(n, seed=n)
    want_cur, want_prev = oracle_run(pos, iters)
    cur, prev = seg_step(hgb, pos, iters, variant, seg=seg, want_prev=True)
    assert np.array_equal(cur, want_cur) and np.array_equal(prev, want_prev)


@pytest.mark.parametrize("variant", [4242, 5242, 6242, 7242])
@pytest.mark.parametrize("seg,rows,weight", [(32, 1, 1), (64, 1, 1), (96, 3, 2), (352, 2, 5), (2048, 16, 16), (96, 256, 1), (0, 0, 0)])
def test_seg_every_schedule_gives_the_same_bits(hgb, variant, seg, rows, weight):
    from oracle import port
    for n in (3000, 777):
        pos = synth.plummer(n, seed=7)
        want = np.zeros_like(pos)
        port.nbody_step_range(pos, want, 0, n)
        got = seg_step(hgb, pos, 1, variant, seg, rows, weight)
        assert np.array_equal(got, want), (variant, seg, rows, weight, n)


@pytest.mark.parametrize("name,gen,n,iters", [("uniform", synth.uniform_cube, 2049, 10),
                                              ("plummer", synth.plummer, 4096, 10)])
def test_seg_reference_golden_10_steps(hgb, golden, name, gen, n, iters):
    g = golden["nbody"][f"{name}_n{n}_it{iters}"]
    cur, prev = seg_step(hgb, gen(n), iters, seg=256, want_prev=True)
    assert sha(cur) == g["planets_sha256"] and sha(prev) == g["buffer_sha256"]


def test_seg_out_of_range_operands_and_duplicates(hgb):
    from oracle import port
    n = 2500
    pos = synth.uniform_cube(n, seed=3)
    pos[17] = [1e120, -3e150, 2e103]
    pos[1999] = [4e-200, 1e-310, -2e-308]
    pos[100:200] = pos[300:400]
    for poison in (False, True):
        if poison:
            pos[1234, 1] = np.nan
        want = np.zeros_like(pos)
        port.nbody_step_range(pos, want, 0, n)
        got = seg_step(hgb, pos, 1, seg=128)
        assert np.array_equal(np.isnan(got), np.isnan(want))
        ok = ~np.isnan(want)
        assert np.array_equal(got[ok], want[ok])


def test_seg_many_steps_on_one_handle(hgb):
    n = 700
    pos = synth.uniform_cube(n, seed=21)
    want4, _ = oracle_run(pos, 4)
    with hgb.nbody.NBody(n) as nb:
        nb.symmetric(4242, 0).segment(64, 2, 3).upload(pos)
        nb.step(37)
        nb.upload(pos)
        nb.step(4)
        assert np.array_equal(nb.download(), want4)


def test_seg_64k_plummer_all_bodies_equal_the_standard_kernel(hgb):
    from oracle import port
    n = 65536
    pos = synth.plummer(n, seed=1234)
    targets = (synth.splitmix64(99, 512) % np.uint64(n)).astype(np.int32)
    want = port.nbody_accel_subset(pos, targets)
    with hgb.nbody.NBody(n) as nb:
        nb.symmetric(0, 0).upload(pos)
        nb.step(2)
        std = nb.download()
    assert np.array_equal(seg_step(hgb, pos, 1)[targets], want)
    assert np.array_equal(seg_step(hgb, pos, 2), std)
    assert np.array_equal(seg_step(hgb, pos, 2, 6242), std)


def test_seg_full_size_1m_all_bodies_equal_the_standard_kernel(hgb):
    from oracle import port
    n = 1 << 20
    pos = synth.uniform_cube(n, seed=1)
    targets = (synth.splitmix64(5, 1024) % np.uint64(n)).astype(np.int32)
    want = port.nbody_accel_subset(pos, targets)
    with hgb.nbody.NBody(n) as nb:
        nb.symmetric(0, 0).upload(pos)
        nb.step(1)
        std = nb.download()
    for variant in (4242, 6242):
        got = seg_step(hgb, pos, 1, variant)
        assert np.array_equal(got[targets], want)
        assert np.array_equal(got, std)


def _gpu_counts(hgb, monkeypatch):
    g = min(hgb.lib().hgb_device_count(), 8)
    if g < 2:
        monkeypatch.setenv("HGB_VIRTUAL_GPUS", "4")
        return [2, 3, 4]
    return sorted({2, g})


@pytest.mark.parametrize("n,iters,seg", [(5000, 3, 64), (2048, 2, 32), (70001, 2, 1024), (100, 2, 32), (40000, 1, 0)])
def test_seg_blocks_dealt_over_gpus_are_bit_identical(hgb, monkeypatch, n, iters, seg):
    """blocks of 64 targets dealt over the GPUs in turns (peer-mapped chain words, polling finish kernel): same bits"""
    from oracle import port
    counts = _gpu_counts(hgb, monkeypatch)
    pos = synth.plummer(n, seed=n)
    one_cur, one_prev = seg_step(hgb, pos, iters, 6242, seg=seg, want_prev=True)
    if n <= 5000:
        want_cur, want_prev = oracle_run(pos, iters)
        assert np.array_equal(one_cur, want_cur) and np.array_equal(one_prev, want_prev)
    sample = np.random.default_rng(n).choice(n, size=min(n, 96), replace=False).astype(np.int32)
    assert np.array_equal(one_cur[sample], port.nbody_accel_subset(one_prev, sample))
    for g in counts:
        for variant, rows, deal in ((6242, 0, 0), (6242, 1, 1), (4242, 2, 6), (7242, 3, 3)):
            cur, prev = seg_step(hgb, pos, iters, variant, seg=seg, rows=rows, weight=max(1, rows), want_prev=True, ngpus=g, deal=deal)
            assert np.array_equal(cur, one_cur) and np.array_equal(prev, one_prev), (g, variant, rows, deal)


def test_seg_abandoned_step_is_reported_even_when_more_steps_follow(hgb, monkeypatch):
    """A hand-off that never arrives makes the kernel give up and raise its abort flag; the host reads the flag
    once, after the LAST step of an advance.  The flag therefore has to survive the per-step reset of the control
    words of the steps in between (it did not: a full memset before every step erased it and the run returned
    HGB_OK on garbage).  HGB_NBODY_SEG_INJECT_ABORT=k (read when the handle is made) raises the flag before
    one-GPU segmented step k, which is how a lost hand-off looks to everything downstream."""
    n = 3000
    pos = synth.plummer(n, seed=11)
    monkeypatch.setenv("HGB_NBODY_SEG_INJECT_ABORT", "0")
    with hgb.nbody.NBody(n) as nb:
        monkeypatch.delenv("HGB_NBODY_SEG_INJECT_ABORT")
        nb.symmetric(6242, 0).upload(pos)
        with pytest.raises(hgb.HgbError) as e:
            nb.step(3)                                  # abandoned in step 0, two more steps after it
        assert e.value.code == hgb._abi.HGB_ERR_STATE and "abandoned" in str(e.value)
        # reported once; the same handle then works again (the injection was for step 0 only)
        nb.upload(pos)
        nb.step(2)
        got = nb.download()
    want, _ = oracle_run(pos, 2)
    assert np.array_equal(got, want)
