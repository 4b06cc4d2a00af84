"""Host-side mirror of the reference's nbody interface (src/nbody/nbody-run.h).

`run` keeps the reference's name, argument order and in-place semantics;
`NBody` is the device-resident handle (upload once, step many times).
"""
import ctypes as C

import numpy as np

from . import _abi
from ._abi import check, dptr, lib

G = 9.8      # reference src/nbody/nbody-run.h:12
EPS = 0.005  # reference src/nbody/nbody-run.h:13


def _bodies(a, name):
    if not (isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.c_contiguous
            and a.ndim == 2 and a.shape[1] == 3):
        raise TypeError(f"{name}: expected a C-contiguous float64 array of shape (n, 3) "
                        "(== Float3D[n], reference src/nbody/nbody-util.h:19-23)")
    return a


def run(planets, buffer, number_of_planets, iterations, number_of_processes=1):
    """void run(Float3D*, Float3D*, int, int, int)  -- reference src/nbody/nbody-run.c:27-37.

    In place, on the GPU.  As in the reference the two LOCAL pointers are swapped, so
    state k ends in `planets` when k is even and in `buffer` when k is odd.
    `number_of_processes` is the number of GPUs the targets are sharded over.
    """
    _bodies(planets, "planets")
    _bodies(buffer, "buffer")
    if planets.shape[0] < number_of_planets or buffer.shape[0] < number_of_planets:
        raise ValueError("arrays shorter than number_of_planets")
    check(lib().hgb_nbody_run(dptr(planets), dptr(buffer), int(number_of_planets), int(iterations),
                              int(number_of_processes)))


def set_mode(mode):
    """EXACT (bit-identical to the reference, default) or FAST (FMA + rsqrt-Newton)."""
    check(lib().hgb_set_nbody_mode(int(mode)))


def set_exchange(kind):
    check(lib().hgb_set_exchange(int(kind)))


class NBody:
    """Device-resident body set.  ngpus=None -> one process per GPU (uses the rank
    communicator set up by hg_c_benchmarks_b200.dist); ngpus=g -> this process drives g GPUs."""

    def __init__(self, n, ngpus=1, mode=_abi.NBODY_EXACT, exchange=_abi.EXCHANGE_NCCL):
        self.n = int(n)
        self._h = C.c_void_p()
        if ngpus is None:
            check(lib().hgb_nbody_create_rank(C.byref(self._h), self.n))
        else:
            check(lib().hgb_nbody_create(C.byref(self._h), self.n, int(ngpus)))
        check(lib().hgb_nbody_config(self._h, int(mode), int(exchange)))

    def tune(self, block=0, unroll=0, warps=0):
        check(lib().hgb_nbody_tune(self._h, int(block), int(unroll), int(warps)))
        return self

    def symmetric(self, variant=1, min_n=0):
        """Pair-symmetric EXACT step (each pair evaluated once, same per-target summation order,
        bit-identical; hgb_nbody_set_symmetric).  variant 0 = off, 1 = default, else
        R*100 + U*10 + CTAs per SM; applies to one-GPU handles with n >= min_n."""
        check(lib().hgb_nbody_set_symmetric(self._h, int(variant), int(min_n)))
        return self

    def segment(self, seg=0, rows=0, weight=0):
        """schedule of the second-generation pair-symmetric kernel: sources per item (multiple of 32), rows
        per tile, weight of the segment index in the ticket order (0 = default each)"""
        check(lib().hgb_nbody_set_segment(self._h, int(seg), int(rows), int(weight)))
        return self

    def deal(self, rows_per_turn=0):
        """rows of targets per turn of a GPU when the wavefront is dealt over several GPUs (0 = default)"""
        check(lib().hgb_nbody_set_deal(self._h, int(rows_per_turn)))
        return self

    def describe(self):
        """(kernel, parallelism) the next step() will use on this handle as configured now"""
        buf = C.create_string_buffer(512)
        check(lib().hgb_nbody_describe(self._h, buf, 512))
        kernel, _, par = buf.value.decode().partition(" | ")
        return kernel, par

    def upload(self, pos):
        _bodies(pos, "pos")
        assert pos.shape[0] == self.n
        check(lib().hgb_nbody_upload(self._h, dptr(pos)))
        return self

    def step(self, iterations=1):
        """Run `iterations` steps; returns device milliseconds (CUDA events)."""
        ms = C.c_float(0.0)
        check(lib().hgb_nbody_step(self._h, int(iterations), C.byref(ms)))
        return ms.value

    def download(self, want_prev=False):
        cur = np.empty((self.n, 3), dtype=np.float64)
        prev = np.empty((self.n, 3), dtype=np.float64) if want_prev else None
        check(lib().hgb_nbody_download(self._h, dptr(cur), dptr(prev)))
        return (cur, prev) if want_prev else cur

    def accel_subset(self, targets):
        targets = np.ascontiguousarray(targets, dtype=np.int32)
        out = np.empty((targets.size, 3), dtype=np.float64)
        check(lib().hgb_nbody_accel_subset(self._h, targets.ctypes.data_as(C.POINTER(C.c_int)),
                                           targets.size, dptr(out)))
        return out

    def energy(self):
        e = C.c_double(0.0)
        check(lib().hgb_nbody_energy(self._h, C.byref(e)))
        return e.value

    # ---- integrator extension (the reference has positions only) ----
    def set_dynamics(self, vel=None, mass=None, dt=1e-3):
        """velocities (n,3) (None = at rest), masses (n,) (None = 1.0) and the time step of
        integrate(): a = G sum_j m_j d f ; v += dt a ; x += dt v, every op one IEEE rounding."""
        if vel is not None:
            _bodies(vel, "vel")
            assert vel.shape[0] == self.n
        if mass is not None:
            mass = np.ascontiguousarray(mass, dtype=np.float64)
            assert mass.shape == (self.n,)
        check(lib().hgb_nbody_set_dynamics(self._h, dptr(vel), dptr(mass), float(dt)))
        return self

    def integrate(self, steps=1):
        ms = C.c_float(0.0)
        check(lib().hgb_nbody_integrate(self._h, int(steps), C.byref(ms)))
        return ms.value

    def velocities(self):
        vel = np.empty((self.n, 3), dtype=np.float64)
        check(lib().hgb_nbody_download_velocities(self._h, dptr(vel)))
        return vel

    def total_energy(self):
        """(kinetic, potential) of the current state"""
        k, u = C.c_double(0.0), C.c_double(0.0)
        check(lib().hgb_nbody_total_energy(self._h, C.byref(k), C.byref(u)))
        return k.value, u.value

    def close(self):
        if self._h:
            lib().hgb_nbody_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
