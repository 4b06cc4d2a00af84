"""hg_c_benchmarks_b200 -- the hg-c-benchmarks hot path (all-pairs FP64 nbody step,
2D convolution stencil) as hand-written CUDA for B200 (sm_100a) behind a C ABI.

  include/hgb200.h                 the C ABI (libhgb200.so)
  hg_c_benchmarks_b200/csrc/       CUDA kernels + runtime
  hg_c_benchmarks_b200.nbody       mirror of reference src/nbody/nbody-run.h
  hg_c_benchmarks_b200.convolution mirror of reference src/convolution/convolution-{run,util}.h
  hg_c_benchmarks_b200.dist        shard plans + one-process-per-GPU communicator
  hg_c_benchmarks_b200.synth       seeded synthetic inputs

No CPU fallback: compute entry points raise when libhgb200.so or a GPU is missing.
"""
from . import _abi, convolution, dist, nbody, synth  # noqa: F401
from ._abi import (CONV_EXACT, CONV_FAST, EXCHANGE_NCCL, EXCHANGE_P2P, NBODY_EXACT,  # noqa: F401
                   NBODY_FAST, HgbError, lib)

__all__ = ["nbody", "convolution", "dist", "synth", "lib", "HgbError", "NBODY_EXACT",
           "NBODY_FAST", "EXCHANGE_NCCL", "EXCHANGE_P2P", "CONV_EXACT", "CONV_FAST"]
