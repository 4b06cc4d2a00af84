"""GPU tests for the device-resident image lifecycle: add_padding,
copy_padded_image (backup / restore), remove_padding and get_checksum around the benchmark loop
of the reference's main (src/2d-convolution.c:41-74), without host round trips.  Bit-identical
to the host-side versions / the oracle."""
import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("w,h,p", [(1, 1, 0), (1, 1, 1), (7, 5, 3), (300, 70, 1), (513, 260, 2), (257, 33, 6)])
def test_upload_dense_is_add_padding(hgb, w, h, p):
    from oracle import port
    dense = synth.image_float(w, h, seed=w + h)
    with hgb.convolution.Conv2D(w, h, p) as cv:
        cv.upload_dense(dense)
        got = cv.download()
        assert np.array_equal(got, port.add_padding(dense, p))          # pad columns made on the device
        assert np.array_equal(cv.download_dense(), dense)               # remove_padding


def test_repetition_loop_backup_restore_checksum(hgb, shipped_image):
    """the REPETITION loop of src/2d-convolution.c:63-74 kept on the device: restore, run, checksum"""
    from oracle import port
    h, w = shipped_image.shape
    p, it = 2, 10
    want_cur = port.add_padding(shipped_image, p)
    want_prev = want_cur.copy()
    port.conv_run(want_cur, want_prev, w, h, p, synth.KERNEL_TXT, it)
    want_sum = port.conv_checksum(want_cur, w, h, p)
    assert f"{want_sum:f}" == "34779110.000000"                            # BASELINE.md / SURVEY 8c
    with hgb.convolution.Conv2D(w, h, p) as cv:
        cv.set_kernel(synth.KERNEL_TXT).upload_dense(shipped_image).backup()
        for rep in range(3):
            cv.restore()
            assert np.array_equal(cv.download_dense(), shipped_image)
            cv.step(it)
            assert cv.checksum() == want_sum
            assert np.array_equal(cv.download(), want_cur)
            assert np.array_equal(cv.download_dense(), want_cur[:, p:p + w])


def test_backup_is_a_copy_not_an_alias(hgb):
    w, h, p = 130, 40, 1
    dense = synth.image_float(w, h, seed=2)
    other = synth.image_float(w, h, seed=3)
    with hgb.convolution.Conv2D(w, h, p) as cv:
        cv.set_kernel(synth.LAPLACE_TXT)
        with pytest.raises(hgb.HgbError) as e:
            cv.restore()
        assert e.value.code == hgb._abi.HGB_ERR_STATE
        with pytest.raises(hgb.HgbError):
            cv.backup()                                                # nothing uploaded yet
        cv.upload_dense(dense).backup()
        cv.upload_dense(other)                                          # overwrite the live image
        cv.step(3)
        cv.restore()
        assert np.array_equal(cv.download_dense(), dense)


@pytest.mark.parametrize("w,h,k,it", [(300, 70, 3, 4), (513, 1300, 7, 9), (260, 2049, 5, 10)])
def test_lifecycle_over_row_bands(hgb, monkeypatch, w, h, k, it):
    g = min(hgb.lib().hgb_device_count(), 8)
    if g < 2:
        monkeypatch.setenv("HGB_VIRTUAL_GPUS", "4")
        g = 4
    p = k // 2
    dense = synth.image_float(w, h, seed=w)
    kern = synth.kernel_dense(k)
    with hgb.convolution.Conv2D(w, h, p, ngpus=1) as one:
        one.set_kernel(kern).upload_dense(dense)
        one.step(it)
        want, want_sum = one.download(), one.checksum()
    for gg in sorted({2, g}):
        with hgb.convolution.Conv2D(w, h, p, ngpus=gg) as cv:
            cv.set_exchange(hgb.EXCHANGE_P2P).set_kernel(kern).upload_dense(dense).backup()
            for rep in range(2):
                cv.restore()
                cv.step(it)
                assert np.array_equal(cv.download(), want), (gg, rep)
                assert np.array_equal(cv.download_dense(), want[:, p:p + w])
                assert cv.checksum() == want_sum
