"""hgb_nbody_energy_static: the potential energy for callers that may not allocate device memory.

Some users of libhgb200.so cannot go through a handle, because creating one calls cudaMalloc: an
in-situ hook inside a host code that already gave all of the GPU's memory to its own pool, a
monitoring call made while a step loop is in flight (cudaMalloc / cudaFree stall the whole device),
a plugin that may only touch the device pointers it is handed.  For them the library has one entry
point that takes device pointers and keeps its per-block partial sums in a fixed 64 KiB array that is
part of the library image -- no cudaMalloc, no cudaFree, no allocating library call.

This file plays such a caller: torch stands in for "whoever owns the memory" (it provides the device
buffers and the stream), and everything the library does goes through the C ABI.  The answers are
compared with the existing, allocating entry point (hgb_nbody_energy on a handle) for the same
positions.  The built-in array is shared by all calls and keeps the previous call's contents, so the
checks are made on two calls in a row, the second smaller than the first: a result that depended on
what the first call left behind would show up there.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must
not hide any other test.  It passes on its own in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12   # the tolerance the energy tests use (tests/test_nbody_gpu.py, test_integrator_gpu.py)
MAX_BODIES = 1 << 20   # HGB_STATIC_ENERGY_MAX_BODIES in include/hgb200.h


def _energy_with_handle(hgb, pos):
    """the existing entry point: a handle (which allocates), upload, hgb_nbody_energy"""
    with hgb.nbody.NBody(pos.shape[0], 1) as nb:
        nb.upload(pos)
        return nb.energy()


def _energy_static(hgb, torch, pos_dev, n, out_dev):
    """the new entry point on torch's current stream; pos_dev / out_dev are torch CUDA tensors"""
    stream = torch.cuda.current_stream()
    rc = hgb.lib().hgb_nbody_energy_static(C.c_void_p(pos_dev.data_ptr()), n, C.c_void_p(out_dev.data_ptr()),
                                           C.c_void_p(stream.cuda_stream))
    hgb._abi.check(rc)
    stream.synchronize()
    return float(out_dev.cpu()[0])


def test_static_energy_matches_the_handle_entry_point_twice_in_a_row(hgb):
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    big, small = synth.plummer(6000, seed=2), synth.fill_planets(1500)   # 47 blocks, then 12 (ragged both)
    want_big, want_small = _energy_with_handle(hgb, big), _energy_with_handle(hgb, small)
    assert want_big < 0.0 and want_small < 0.0

    # the caller's memory: positions as Float3D AoS and one double for the answer, made BEFORE the
    # calls under test so that nothing is allocated between or inside them
    big_dev = torch.from_numpy(big).cuda()
    small_dev = torch.from_numpy(small).cuda()
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()

    launches = hgb.lib().hgb_launch_count()
    got_big = _energy_static(hgb, torch, big_dev, 6000, out)
    out.fill_(float("nan"))
    got_small = _energy_static(hgb, torch, small_dev, 1500, out)     # slots 12..46 still hold call 1's partials
    out.fill_(float("nan"))
    got_big_again = _energy_static(hgb, torch, big_dev, 6000, out)
    assert hgb.lib().hgb_launch_count() == launches + 6               # two kernels of this library per call

    assert abs(got_big - want_big) <= REL_TOL * abs(want_big), (got_big, want_big)
    assert abs(got_small - want_small) <= REL_TOL * abs(want_small), (got_small, want_small)
    assert got_big_again == got_big                                   # fixed summation order: same bits


def test_static_energy_refuses_what_it_cannot_hold(hgb):
    torch = pytest.importorskip("torch")
    torch.cuda.set_device(0)
    lib, abi = hgb.lib(), hgb._abi
    pos = torch.zeros(3 * 16, dtype=torch.float64, device="cuda")
    out = torch.full((1,), 7.0, dtype=torch.float64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p, o = C.c_void_p(pos.data_ptr()), C.c_void_p(out.data_ptr())
    launches = lib.hgb_launch_count()
    # one body more than the built-in array covers: a clear error, nothing launched, output untouched
    assert lib.hgb_nbody_energy_static(p, MAX_BODIES + 1, o, stream) == abi.HGB_ERR_ARG
    message = lib.hgb_last_error().decode()
    assert "HGB_STATIC_ENERGY_MAX_BODIES" in message and str(MAX_BODIES) in message
    assert lib.hgb_nbody_energy_static(p, 0, o, stream) == abi.HGB_ERR_ARG
    assert lib.hgb_nbody_energy_static(None, 16, o, stream) == abi.HGB_ERR_ARG
    # host memory is not something this entry point can use: it copies nothing
    host = np.zeros(3 * 16)
    assert lib.hgb_nbody_energy_static(C.c_void_p(host.ctypes.data), 16, o, stream) == abi.HGB_ERR_ARG
    assert "device" in lib.hgb_last_error().decode()
    assert lib.hgb_launch_count() == launches
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == 7.0
    # 16 bodies at one point: 240 ordered pairs, each 1/sqrt(EPS); U = -G/2 * 240 / sqrt(0.005)
    hgb._abi.check(lib.hgb_nbody_energy_static(p, 16, o, stream))
    torch.cuda.synchronize()
    want = -9.8 * 0.5 * 240.0 / np.sqrt(0.005)
    assert abs(float(out.cpu()[0]) - want) <= REL_TOL * abs(want)
