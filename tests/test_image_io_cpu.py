"""Image files: the reference's text format read with the reference's token rules
(read_image_from_fd, src/convolution/convolution-util.c:9-45) without its line-length limit, plus the
binary format.  Host code: runs without a GPU.  Checked against the reference's own parser where
oracle/_ref holds it (this container), and against the committed parse of the shipped input."""
import os
import struct

import numpy as np
import pytest

from hg_c_benchmarks_b200 import convolution as cv
from hg_c_benchmarks_b200 import synth
from hg_c_benchmarks_b200._abi import HgbError

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _write_reference_style(path, img):
    h, w = img.shape
    with open(path, "w") as f:
        f.write(f"{h} {w}\n")
        for row in img:
            f.write(" ".join(f"{int(v)}" if v == int(v) else repr(float(v)) for v in row) + "\n")


def test_shipped_input_parses_like_the_reference(tmp_path, shipped_image):
    p = tmp_path / "img.input"
    _write_reference_style(p, shipped_image)
    img = cv.read_image(p)
    assert (img.height, img.width) == shipped_image.shape
    assert np.array_equal(img.image, shipped_image)
    assert cv.get_checksum(img) == float(shipped_image.sum())


def test_height_comes_first_and_kernels_parse(tmp_path):
    p = tmp_path / "k.txt"
    p.write_text("2 3\n0 0.125 -1\n1e-3 .5 7\n")           # 2 rows, 3 columns
    img = cv.read_image(p)
    assert (img.height, img.width) == (2, 3)
    assert np.array_equal(img.image, [[0, 0.125, -1], [1e-3, 0.5, 7]])
    q = tmp_path / "laplace.txt"
    q.write_text("3 3\n0 0.25 0\n0.25 -1 0.25\n0 0.25 0\n")
    assert np.array_equal(cv.read_image(q).image, synth.LAPLACE_TXT)


def test_token_rules_match_strtok_strtod(tmp_path):
    p = tmp_path / "odd.txt"
    # runs of spaces collapse; surplus tokens and lines are ignored; a trailing-space line yields a
    # "\n" token that strtod reads as 0.0; non-numeric text reads as 0.0 -- all as in the reference
    p.write_text("2 3 99\n1   2 \nx 5 6 7 8\nignored line\n")
    img = cv.read_image(p)
    assert np.array_equal(img.image, [[1, 2, 0], [0, 5, 6]])


@pytest.mark.parametrize("text,needle", [("", "empty"), ("3\n", "width"), ("2 2\n1 2\n", "enough lines"),
                                         ("2 2\n1 2\n3\n", "dimensions"), ("0 5\n", "extents")])
def test_malformed_files_are_errors_not_exits(tmp_path, text, needle):
    p = tmp_path / "bad.txt"
    p.write_text(text)
    with pytest.raises(HgbError) as e:
        cv.read_image(p)
    assert needle in str(e.value)
    with pytest.raises(HgbError):
        cv.read_image(tmp_path / "missing.txt")


def test_rows_wider_than_the_reference_line_buffer(tmp_path):
    """a 20000-pixel row is ~380 KB of text: over the reference's 16384-byte fgets buffer"""
    img = cv.Image(20000, 3)
    img.image[...] = synth.image_float(20000, 3, seed=11)
    p = tmp_path / "wide.txt"
    cv.write_image(p, img)
    assert os.path.getsize(p) > 3 * 16384
    back = cv.read_image(p)
    assert np.array_equal(back.image, img.image)              # %.17g round-trips every double


def test_binary_round_trip_and_layout(tmp_path):
    img = cv.Image(37, 5)
    img.image[...] = synth.image_float(37, 5, seed=4) - 0.5
    img.image[0, 0], img.image[1, 1] = -0.0, 5e-324
    p = tmp_path / "img.hgb"
    cv.write_image_binary(p, img)
    raw = p.read_bytes()
    assert raw[:8] == b"HGBIMG1\n" and struct.unpack("<ii", raw[8:16]) == (5, 37)   # height first
    assert len(raw) == 16 + 8 * 37 * 5
    back = cv.read_image_binary(p)
    assert back.image.tobytes() == img.image.tobytes()
    (tmp_path / "short.hgb").write_bytes(raw[:-8])
    with pytest.raises(HgbError):
        cv.read_image_binary(tmp_path / "short.hgb")
    with pytest.raises(HgbError):
        cv.read_image_binary(tmp_path / "k.txt")
    # a crafted header must not size the allocation: huge extents are refused before malloc
    # (2^31-1 squared * 8 overflows size_t arithmetic on the element count times 8)
    for h, w in ((0x7fffffff, 0x7fffffff), (1 << 20, 1 << 20), (5, 38)):
        (tmp_path / "evil.hgb").write_bytes(b"HGBIMG1\n" + struct.pack("<ii", h, w) + raw[16:])
        with pytest.raises(HgbError):
            cv.read_image_binary(tmp_path / "evil.hgb")


def test_against_the_reference_parser_when_built(tmp_path):
    """oracle/_ref/libhgref_conv.so is the unmodified reference: parse the same files with both"""
    so = os.path.join(ROOT, "oracle", "_ref", "libhgref_conv.so")
    if not os.path.exists(so):
        pytest.skip("reference not built here")
    from oracle import ref
    for name, text in [("a.txt", "2 3\n1 2 3\n4 5 6\n"), ("b.txt", "2 3 99\n1   2 \nx 5 6 7 8\nzzz\n"),
                       ("c.txt", "3 1\n-1.5e3\n0x10\n  7\n")]:
        p = tmp_path / name
        p.write_text(text)
        want = ref.read_image(str(p))
        got = cv.read_image(p)
        assert np.array_equal(got.image, want), name
