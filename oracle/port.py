"""ctypes loader for oracle/libhgoracle.so (our C restatement).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def build():
    subprocess.run(["make", "-C", _HERE, "libhgoracle.so"], check=True, capture_output=True)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libhgoracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.oracle_nbody_run.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int]
        L.oracle_nbody_step_range.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int, C.c_int]
        L.oracle_nbody_accel_subset.argtypes = [_dp, C.c_int, _ip, C.c_int, _dp, C.c_int]
        L.oracle_conv_run.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int, _dp, C.c_int, C.c_int,
                                      C.c_int, C.c_int]
        L.oracle_conv_step_rows.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int, _dp, C.c_int,
                                            C.c_int, C.c_int, C.c_int, C.c_int]
        L.oracle_conv_checksum.argtypes = [_dp, C.c_int, C.c_int, C.c_int]
        L.oracle_conv_checksum.restype = C.c_double
        L.oracle_add_padding.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int]
        L.oracle_fill_planets.argtypes = [_dp, C.c_int]
        L.oracle_nbody_energy.argtypes = [_dp, C.c_int, C.c_int]
        L.oracle_nbody_energy.restype = C.c_double
        L.oracle_nbody_integrate.argtypes = [_dp, _dp, _dp, C.c_int, C.c_double, C.c_int, C.c_int, _dp]
        L.oracle_nbody_total_energy.argtypes = [_dp, _dp, _dp, C.c_int, C.c_int, _dp, _dp]
        _LIB = L
    return _LIB


def _p(a):
    assert a.dtype == np.float64 and a.flags.c_contiguous
    return a.ctypes.data_as(_dp)


def ncpu():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def nbody_run(planets, buffer, iterations, threads=None):
    """In place, reference run() semantics (odd/even ping-pong parity)."""
    n = planets.shape[0]
    lib().oracle_nbody_run(_p(planets), _p(buffer), n, iterations, threads or ncpu())


def nbody_step_range(pos, out, t0, t1, threads=None):
    lib().oracle_nbody_step_range(_p(pos), _p(out), pos.shape[0], t0, t1, threads or ncpu())


def nbody_accel_subset(pos, targets, threads=None):
    targets = np.ascontiguousarray(targets, dtype=np.int32)
    out = np.empty((targets.size, 3), dtype=np.float64)
    lib().oracle_nbody_accel_subset(_p(pos), pos.shape[0], targets.ctypes.data_as(_ip),
                                    targets.size, _p(out), threads or ncpu())
    return out


def conv_run(cur, prev, w, h, p, kernel, iterations, threads=None):
    """cur/prev: (h, w+2p) float64 in place.  cur <- state n, prev <- state n-1."""
    kernel = np.ascontiguousarray(kernel, dtype=np.float64)
    kh, kw = kernel.shape
    lib().oracle_conv_run(_p(cur), _p(prev), w, h, p, _p(kernel), kw, kh, iterations,
                          threads or ncpu())


def conv_step_rows(inp, out, w, h, p, kernel, r0, r1, threads=None):
    kernel = np.ascontiguousarray(kernel, dtype=np.float64)
    kh, kw = kernel.shape
    lib().oracle_conv_step_rows(_p(inp), _p(out), w, h, p, _p(kernel), kw, kh, r0, r1,
                                threads or ncpu())


def conv_checksum(flat, w, h, p):
    return lib().oracle_conv_checksum(_p(flat), w, h, p)


def add_padding(dense, p):
    dense = np.ascontiguousarray(dense, dtype=np.float64)
    h, w = dense.shape
    flat = np.empty((h, w + 2 * p), dtype=np.float64)
    lib().oracle_add_padding(_p(dense), _p(flat), w, h, p)
    return flat


def fill_planets(n):
    pos = np.empty((n, 3), dtype=np.float64)
    lib().oracle_fill_planets(_p(pos), n)
    return pos


def nbody_energy(pos, threads=None):
    return lib().oracle_nbody_energy(_p(pos), pos.shape[0], threads or ncpu())


def nbody_integrate(pos, vel, mass, dt, steps, threads=None):
    """Integrator extension (no reference counterpart): pos, vel advanced in place."""
    scratch = np.empty_like(pos)
    lib().oracle_nbody_integrate(_p(pos), _p(vel), _p(mass), pos.shape[0], float(dt), int(steps),
                                 threads or ncpu(), _p(scratch))


def nbody_total_energy(pos, vel, mass, threads=None):
    k, u = C.c_double(0.0), C.c_double(0.0)
    lib().oracle_nbody_total_energy(_p(pos), _p(vel), _p(mass), pos.shape[0], threads or ncpu(),
                                    C.byref(k), C.byref(u))
    return k.value, u.value
