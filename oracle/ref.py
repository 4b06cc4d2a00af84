"""ctypes loader for oracle/_ref/ (the UNMODIFIED reference, compiled by oracle/Makefile).

TEST INFRASTRUCTURE ONLY.  The libraries are prebuilt in the development container
(where /root/reference exists) and travel to the GPU box as binaries; nothing here
reads /root/reference at run time.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF = os.path.join(_HERE, "_ref")
_dp = C.POINTER(C.c_double)
_NB = {}
_CV = {}
# "" = the shipped effective flags (nbody -O2, convolution -O3); "_pgo" = the same sources at
# -O3 with gcc profile feedback, i.e. what the reference's pgo.sh intended (oracle/Makefile)
_VARIANT = ""


def build(reference="/root/reference", pgo=True):
    if not os.path.isdir(reference):
        raise RuntimeError("reference sources not present; oracle/_ref cannot be (re)built here")
    subprocess.run(["make", "-C", _HERE, "ref", "ref-v4"] + (["ref-pgo"] if pgo else []) + [f"REF={reference}"],
                   check=True, capture_output=True)


def _cpu_has_avx512():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    flags = set(line.split(":", 1)[1].split())
                    return {"avx512f", "avx512bw", "avx512cd", "avx512dq", "avx512vl"} <= flags
    except OSError:
        pass
    return False


def available(variant=""):
    """variant "" (x86-64-v3, shipped -O2/-O3), "_pgo" (-O3 + profile feedback) or "_v4" (AVX-512: what
    the reference's -march=native gives on such a host; only offered when the CPU can run it)"""
    if variant == "_v4" and not _cpu_has_avx512():
        return False
    return (os.path.exists(os.path.join(_REF, f"libhgref_nbody{variant}.so"))
            and os.path.exists(os.path.join(_REF, f"libhgref_conv{variant}.so")))


VARIANTS = {"": "shipped flags, -march=x86-64-v3 (-std=c11 -O2 nbody / -O3 convolution, -fopenmp)",
            "_pgo": "-O3 + gcc profile feedback (pgo.sh's intent), -march=x86-64-v3",
            "_v4": "shipped flags, -march=x86-64-v4 (AVX-512: the reference's -march=native on this host)"}


def use(variant=""):
    """select which build of the unmodified reference the functions below call"""
    global _VARIANT
    assert variant in VARIANTS and available(variant), variant
    _VARIANT = variant


def _nb():
    if _VARIANT not in _NB:
        L = C.CDLL(os.path.join(_REF, f"libhgref_nbody{_VARIANT}.so"))
        # void run(Float3D*, Float3D*, int, int, int)   src/nbody/nbody-run.h:46
        L.run.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int]
        L.run.restype = None
        L.fill_planet.argtypes = [_dp, C.c_int]  # src/nbody/nbody-util.h:54
        L.fill_planet.restype = None
        _NB[_VARIANT] = L
    return _NB[_VARIANT]


def _cv():
    if _VARIANT not in _CV:
        L = C.CDLL(os.path.join(_REF, f"libhgref_conv{_VARIANT}.so"))
        L.ref_conv_run_flat.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int, _dp, C.c_int, C.c_int,
                                        C.c_int, C.c_int]
        L.ref_add_padding_flat.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int]
        L.ref_checksum_flat.argtypes = [_dp, C.c_int, C.c_int, C.c_int]
        L.ref_checksum_flat.restype = C.c_double
        L.ref_read_image.argtypes = [C.c_char_p, _dp, C.c_int, C.POINTER(C.c_int),
                                     C.POINTER(C.c_int)]
        L.ref_default_kernel.argtypes = [_dp]
        _CV[_VARIANT] = L
    return _CV[_VARIANT]


def _p(a):
    assert a.dtype == np.float64 and a.flags.c_contiguous
    return a.ctypes.data_as(_dp)


def nbody_run(planets, buffer, iterations, procs=1):
    _nb().run(_p(planets), _p(buffer), planets.shape[0], iterations, procs)


def fill_planets(n):
    pos = np.empty((n, 3), dtype=np.float64)
    for i in range(n):
        _nb().fill_planet(_p(pos[i:i + 1]), i)
    return pos


def conv_run(cur, prev, w, h, p, kernel, iterations, procs=1):
    kernel = np.ascontiguousarray(kernel, dtype=np.float64)
    kh, kw = kernel.shape
    rc = _cv().ref_conv_run_flat(_p(cur), _p(prev), w, h, p, _p(kernel), kw, kh, iterations, procs)
    assert rc == 0


def add_padding(dense, p):
    dense = np.ascontiguousarray(dense, dtype=np.float64)
    h, w = dense.shape
    flat = np.empty((h, w + 2 * p), dtype=np.float64)
    rc = _cv().ref_add_padding_flat(_p(dense), _p(flat), w, h, p)
    assert rc == 0, "reference pad rows are not aliases of the edge rows?"
    return flat


def conv_checksum(flat, w, h, p):
    return _cv().ref_checksum_flat(_p(flat), w, h, p)


def read_image(path, cap=1 << 22):
    out = np.empty(cap, dtype=np.float64)
    w, h = C.c_int(), C.c_int()
    rc = _cv().ref_read_image(path.encode(), _p(out), cap, C.byref(w), C.byref(h))
    assert rc == 0, rc
    return out[: w.value * h.value].reshape(h.value, w.value).copy()


def default_kernel():
    k = np.empty(25, dtype=np.float64)
    _cv().ref_default_kernel(_p(k))
    return k.reshape(5, 5)


def cli(binary, args, timeout=600):
    """Run a shipped CLI driver (writes ../*.res relative to its cwd)."""
    cwd = os.path.join(_REF, "cwd")
    os.makedirs(cwd, exist_ok=True)
    return subprocess.run([os.path.join(_REF, binary)] + [str(a) for a in args], cwd=cwd,
                          capture_output=True, text=True, timeout=timeout, check=True)
