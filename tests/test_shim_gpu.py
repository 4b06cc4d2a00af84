"""The reference's OWN drivers and gtest suites, unmodified, linked against the GPU shim
(shim/Makefile -> oracle/_ref/*_b200, built where /root/reference exists and shipped as
binaries).  This is the north-star's "existing gtest suites drive the GPU path"."""
import os
import subprocess

import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
CWD = os.path.join(REF, "cwd")


def _need(name):
    path = os.path.join(REF, name)
    if not os.path.exists(path):
        pytest.skip(f"{name} not built (needs the reference tree at build time)")
    os.makedirs(CWD, exist_ok=True)
    return path


def test_reference_convolution_gtest_suite_on_gpu():
    exe = _need("convolution_tests_b200")
    r = subprocess.run([exe], cwd=CWD, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "[       OK ] run_on_image.run_laplace_neutralizes_image" in r.stdout
    assert "[  PASSED  ] 11 tests." in r.stdout


def test_reference_nbody_gtest_suite():
    exe = _need("nbody_tests_b200")
    r = subprocess.run([exe], cwd=CWD, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "[  PASSED  ] 3 tests." in r.stdout


@pytest.mark.parametrize("exe_name", ["nbody_symbols_b200", "conv_symbols_b200"])
def test_every_declared_symbol_matches_the_reference_object_code(exe_name):
    """pair_wise_accel / accel / run and apply_kernel_to_padded_point / apply_kernel_to_padded_image /
    run_on_padded_image (reference src/nbody/nbody-run.h:23,34,46, src/convolution/convolution-run.h:
    45-46,58-60,70-71): the shim's symbol and the reference's own compiled function (renamed ref_*)
    in one binary, bit for bit (shim/test/symbols_*.c)."""
    exe = _need(exe_name)
    r = subprocess.run([exe], cwd=CWD, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "PASSED" in r.stdout and " 0 mismatches" in r.stdout


@pytest.mark.parametrize("args", [["-s", "3", "-n", "10"], ["-s", "3", "-n", "2"], ["-s", "3", "-n", "3"]])
def test_reference_nbody_main_on_gpu(golden, args):
    """src/nbody.c as shipped (defaults: 3 bodies, 10 iterations, 10 repetitions) -> ../nbody.res"""
    exe = _need("Nbody-B200")
    r = subprocess.run([exe] + args, cwd=CWD, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.count("Kernel time:") == 10
    got = open(os.path.join(REF, "nbody.res")).read().splitlines()
    assert got == golden["cli"]["nbody " + " ".join(args)]


def _write_image(path, img):
    with open(path, "w") as f:
        f.write(f"{img.shape[0]} {img.shape[1]}\n")       # height then width, as the reader expects
        for row in img:
            f.write(" ".join(repr(float(v)) if v != int(v) else str(int(v)) for v in row) + "\n")


def test_reference_convolution_main_on_gpu(golden, shipped_image, tmp_path):
    """src/2d-convolution.c as shipped on `2d-convolution.input` x `kernel.txt`, n = 10"""
    exe = _need("2D-Convolution-B200")
    img_path, k_path = str(tmp_path / "image.txt"), str(tmp_path / "kernel.txt")
    _write_image(img_path, shipped_image)
    _write_image(k_path, synth.KERNEL_TXT)
    r = subprocess.run([exe, "-f", img_path, "-k", k_path, "-n", "10"], cwd=CWD, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = open(os.path.join(REF, "2d-convolution.res")).read().splitlines()
    assert got == golden["cli"]["conv input kernel.txt n10"]
    assert got == ["34779110.000000"] * 10


def test_reference_main_uses_p_as_gpu_count(golden):
    import hg_c_benchmarks_b200 as H
    if H.lib().hgb_device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    exe = _need("Nbody-B200")
    r = subprocess.run([exe, "-s", "3", "-n", "10", "-p", "2"], cwd=CWD, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert open(os.path.join(REF, "nbody.res")).read().splitlines() == golden["cli"]["nbody -s 3 -n 10"]
