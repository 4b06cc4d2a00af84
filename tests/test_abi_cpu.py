"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports exactly the
symbols include/hgb200.h declares, fails loudly without a GPU (no CPU fallback), and the
Python mirror of the reference's host utilities reproduces the reference's gtest vectors."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "hgb200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hgb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_all_exported_and_bound(hgb):
    names = declared_symbols()
    assert len(names) >= 40
    L = C.CDLL(hgb._abi.LIB_PATH)
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/hgb200.h but not exported"
    assert sorted(hgb._abi.SIGNATURES) == names, "ctypes binding and header disagree"


def test_exported_symbols_are_plain_c(hgb):
    out = subprocess.run(["nm", "-D", "--defined-only", hgb._abi.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    for n in declared_symbols():
        assert n in exported, f"{n} is not an unmangled T symbol"


def test_library_does_not_link_the_oracle(hgb):
    out = subprocess.run(["ldd", hgb._abi.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "hgref" not in out
    # and no product module imports the checker
    pkg = os.path.join(ROOT, "hg_c_benchmarks_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            text = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), fn


def _has_gpu(hgb):
    return hgb.lib().hgb_device_count() > 0


def test_no_cpu_fallback(hgb):
    if _has_gpu(hgb):
        pytest.skip("GPU present")
    a, b = np.zeros((4, 3)), np.zeros((4, 3))
    with pytest.raises(hgb.HgbError) as e:
        hgb.nbody.run(a, b, 4, 2, 1)
    assert e.value.code == hgb._abi.HGB_ERR_NOGPU and "no CPU fallback" in str(e.value)
    cv = hgb.convolution
    img = [cv.add_padding(cv.init_image(4, 4, 1.0), 1)]
    buf = [cv.init_padded_image(4, 4, 1)]
    with pytest.raises(hgb.HgbError):
        cv.run_on_padded_image(img, cv.get_default_kernel(), cv.Args(1, 1, 4, 4), buf)
    with pytest.raises(hgb.HgbError):
        hgb.nbody.NBody(16)
    assert hgb.lib().hgb_launch_count() == 0


def test_missing_library_raises(monkeypatch, hgb):
    monkeypatch.setattr(hgb._abi, "_LIB", None)
    monkeypatch.setattr(hgb._abi, "LIB_PATH", "/nonexistent/libhgb200.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        hgb._abi.lib()


def test_argument_type_checks(hgb):
    with pytest.raises(TypeError):
        hgb.nbody.run(np.zeros((3, 3), dtype=np.float32), np.zeros((3, 3)), 3, 1, 1)
    with pytest.raises(TypeError):
        hgb.nbody.run(np.zeros((3, 4)), np.zeros((3, 4)), 3, 1, 1)
    with pytest.raises(ValueError):
        hgb.nbody.run(np.zeros((2, 3)), np.zeros((2, 3)), 3, 1, 1)


# ---- host mirror of convolution-util: the reference's gtest cases -----------------------
def test_init_image(hgb):
    cv = hgb.convolution
    img = cv.init_image(5, 5, -5)                    # test/convolutionUtilTests.cpp:27-46
    assert img.width == 5 and img.height == 5 and np.all(img.image == -5)
    assert cv.get_checksum(img) == -5 * 5 * 5        # :176-182


def test_init_padded_image(hgb):
    p = hgb.convolution.init_padded_image(3, 3, 2)   # :48-68
    assert (p.width, p.height, p.inner_width, p.inner_height) == (7, 7, 3, 3)
    assert np.all(p.as_matrix() == 0)


def _img3(cv):
    img = cv.init_image(3, 3, 1)
    img.image[...] = [[-1.5, 0, -1], [9, -2, 3], [2, 0, 5]]
    return img


PADDED7 = np.array([[-1.5, -1.5, -1.5, 0, -1, -1, -1]] * 3 + [[9, 9, 9, -2, 3, 3, 3]] +
                   [[2, 2, 2, 0, 5, 5, 5]] * 3, dtype=np.float64)


def test_add_and_remove_padding(hgb):
    cv = hgb.convolution
    p = cv.add_padding(_img3(cv), 2)                 # :96-134
    assert np.array_equal(p.as_matrix(), PADDED7)
    back = cv.remove_padding(p)                      # :136-174
    assert np.array_equal(back.image, _img3(cv).image)
    assert np.all(cv.add_padding(cv.init_image(3, 3, 1), 2).as_matrix() == 1)   # :70-94


def test_update_borders_with_aliased_rows(hgb):
    cv = hgb.convolution                             # :184-231
    p = cv.init_padded_image(3, 3, 2)
    for x in range(7):
        for y in range(7):
            p.set_field(x, y, PADDED7[y][x])
    for (x, y, v) in [(0, 0, 2), (1, 0, 0), (2, 0, 100), (0, 1, 1), (1, 1, 0), (1, 2, -1), (2, 2, 5)]:
        p.set_access(x, y, v)
    cv.update_borders(p)
    want = np.array([[2, 2, 2, 0, 100, 100, 100]] * 3 + [[1, 1, 1, 0, 3, 3, 3]] +
                    [[2, 2, 2, -1, 5, 5, 5]] * 3, dtype=np.float64)
    assert np.array_equal(p.as_matrix(), want)


def test_copy_padded_image_error_codes(hgb):
    cv = hgb.convolution                             # src/convolution/convolution-util.c:357-373
    a, b, c = cv.init_padded_image(3, 3, 1), cv.init_padded_image(3, 3, 1), cv.init_padded_image(4, 3, 1)
    assert cv.copy_padded_image(a, b) == 0
    assert cv.copy_padded_image(None, b) == -1 and cv.copy_padded_image(a, None) == -1
    assert cv.copy_padded_image(a, c) == -2


def test_default_kernel(hgb):
    k = hgb.convolution.get_default_kernel().image
    assert k[2, 2] == -1 and k[0, 0] == 2 and k[0, 1] == -2 and k.sum() == 2 * 13 - 2 * 12 - 3


def test_synthetic_generators_are_stable(hgb):
    s = hgb.synth
    assert s.uniform_cube(4)[0, 0].hex() == s.uniform01(1, 1)[0].hex()
    assert np.array_equal(s.uniform01(9, 10)[3:7], s.uniform01(9, 4, offset=3))
    p = s.plummer(1000)
    assert p.shape == (1000, 3) and np.isfinite(p).all() and np.abs(p).max() < 200
    im = s.image_int(64, 32)
    assert im.min() >= 0 and im.max() <= 255 and np.array_equal(im, np.round(im))
