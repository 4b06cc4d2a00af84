"""ctypes binding of libhgb200.so (the C ABI declared in include/hgb200.h).

There is no CPU fallback anywhere in this package: if the shared library is
missing, or no B200 is visible, every compute entry point raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhgb200.so")

HGB_OK = 0
HGB_ERR_ARG, HGB_ERR_CUDA, HGB_ERR_NCCL, HGB_ERR_STATE, HGB_ERR_NOGPU = -1, -2, -3, -4, -5
NBODY_EXACT, NBODY_FAST = 0, 1
CONV_EXACT, CONV_FAST = 0, 1
EXCHANGE_NCCL, EXCHANGE_P2P = 0, 1
UNIQUE_ID_BYTES = 128

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_fp = C.POINTER(C.c_float)
_vp = C.c_void_p

# name -> (restype, argtypes); must list EVERY symbol include/hgb200.h declares
SIGNATURES = {
    "hgb_version": (C.c_char_p, []),
    "hgb_last_error": (C.c_char_p, []),
    "hgb_device_count": (C.c_int, []),
    "hgb_init": (C.c_int, []),
    "hgb_finalize": (None, []),
    "hgb_launch_count": (C.c_ulonglong, []),
    "hgb_flush_l2": (C.c_int, []),
    "hgb_measure_fp64_peak": (C.c_int, [_dp, _dp]),
    "hgb_measure_hbm_copy": (C.c_int, [_dp]),
    "hgb_comm_unique_id": (C.c_int, [_vp, C.c_int]),
    "hgb_comm_init_rank": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int]),
    "hgb_comm_rank": (C.c_int, []),
    "hgb_comm_world": (C.c_int, []),
    "hgb_comm_barrier": (C.c_int, []),
    "hgb_comm_destroy": (None, []),
    "hgb_nbody_run": (C.c_int, [_dp, _dp, C.c_int, C.c_int, C.c_int]),
    "hgb_set_nbody_mode": (C.c_int, [C.c_int]),
    "hgb_set_exchange": (C.c_int, [C.c_int]),
    "hgb_nbody_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int]),
    "hgb_nbody_create_rank": (C.c_int, [C.POINTER(_vp), C.c_int]),
    "hgb_nbody_config": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hgb_nbody_tune": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int]),
    "hgb_nbody_set_symmetric": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hgb_nbody_symmetric": (C.c_int, [_vp]),
    "hgb_nbody_set_segment": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int]),
    "hgb_nbody_set_deal": (C.c_int, [_vp, C.c_int]),
    "hgb_nbody_describe": (C.c_int, [_vp, C.c_char_p, C.c_int]),
    "hgb_nbody_fp64_instr": (C.c_double, [_vp]),
    "hgb_nbody_sym_counters": (C.c_int, [_vp, C.POINTER(C.c_ulonglong)]),
    "hgb_nbody_upload": (C.c_int, [_vp, _dp]),
    "hgb_nbody_step": (C.c_int, [_vp, C.c_int, _fp]),
    "hgb_nbody_download": (C.c_int, [_vp, _dp, _dp]),
    "hgb_nbody_accel_subset": (C.c_int, [_vp, _ip, C.c_int, _dp]),
    "hgb_nbody_energy": (C.c_int, [_vp, _dp]),
    "hgb_nbody_energy_static": (C.c_int, [_vp, C.c_int, _vp, _vp]),   # device pointers + cudaStream_t
    "hgb_nbody_set_dynamics": (C.c_int, [_vp, _dp, _dp, C.c_double]),
    "hgb_nbody_integrate": (C.c_int, [_vp, C.c_int, _fp]),
    "hgb_nbody_download_velocities": (C.c_int, [_vp, _dp]),
    "hgb_nbody_total_energy": (C.c_int, [_vp, _dp, _dp]),
    "hgb_nbody_destroy": (C.c_int, [_vp]),
    "hgb_nbody_pairwise": (C.c_int, [_dp, _dp, C.c_int, _dp]),
    "hgb_selftest_inv_sqrt": (C.c_int, [C.c_ulonglong, C.c_ulonglong, C.POINTER(C.c_ulonglong), _dp]),
    "hgb_conv_run": (C.c_int, [_dp, _dp, C.c_int, C.c_int, C.c_int, _dp, C.c_int, C.c_int, C.c_int,
                               C.c_int]),
    "hgb_set_conv_mode": (C.c_int, [C.c_int]),
    "hgb_set_conv_streaming": (C.c_int, [C.c_int]),
    "hgb_conv_unpin": (C.c_int, []),
    "hgb_conv_set_mode": (C.c_int, [_vp, C.c_int]),
    "hgb_conv_mode_in_effect": (C.c_int, [_vp]),
    "hgb_conv_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int, C.c_int, C.c_int]),
    "hgb_conv_create_rank": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int, C.c_int]),
    "hgb_conv_set_kernel": (C.c_int, [_vp, _dp, C.c_int, C.c_int]),
    "hgb_conv_tune": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hgb_conv_set_depth": (C.c_int, [_vp, C.c_int]),
    "hgb_conv_set_exchange": (C.c_int, [_vp, C.c_int]),
    "hgb_conv_band": (C.c_int, [_vp, _ip, _ip]),
    "hgb_conv_upload": (C.c_int, [_vp, _dp]),
    "hgb_conv_step": (C.c_int, [_vp, C.c_int, _fp]),
    "hgb_conv_download": (C.c_int, [_vp, _dp, _dp]),
    "hgb_conv_upload_dense": (C.c_int, [_vp, _dp]),
    "hgb_conv_download_dense": (C.c_int, [_vp, _dp]),
    "hgb_conv_backup": (C.c_int, [_vp]),
    "hgb_conv_restore": (C.c_int, [_vp]),
    "hgb_conv_checksum": (C.c_int, [_vp, _dp]),
    "hgb_conv_destroy": (C.c_int, [_vp]),
    "hgb_image_read_text": (C.c_int, [C.c_char_p, _ip, _ip, C.POINTER(_dp)]),
    "hgb_image_write_text": (C.c_int, [C.c_char_p, _dp, C.c_int, C.c_int]),
    "hgb_image_read_binary": (C.c_int, [C.c_char_p, _ip, _ip, C.POINTER(_dp)]),
    "hgb_image_write_binary": (C.c_int, [C.c_char_p, _dp, C.c_int, C.c_int]),
    "hgb_image_free": (None, [_dp]),
}

_LIB = None


class HgbError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"libhgb200 error {code}: {message}")
        self.code = code


def lib():
    """Load libhgb200.so; raises (never falls back) when it has not been built."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (nvcc, sm_100a).  There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)      # AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def check(rc):
    if rc != HGB_OK:
        raise HgbError(rc, lib().hgb_last_error().decode(errors="replace"))
    return rc


def dptr(a):
    import numpy as np
    if a is None:
        return None
    if not (isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.c_contiguous):
        raise TypeError("expected a C-contiguous float64 numpy array (the reference's double data)")
    return a.ctypes.data_as(_dp)
