"""Where the first-generation wavefront waits (diagnostic variant 1242): python tools/sym_hist.py [n]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["HGB_NBODY_SYM_HIST"] = "1"
import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
pos = synth.uniform_cube(n, seed=1)
with H.nbody.NBody(n) as nb:
    nb.symmetric(1242, 0).upload(pos)
    nb.step(1)
    nb.upload(pos)
    ms = nb.step(1)
    c = (C.c_ulonglong * 7)()
    H._abi.check(H.lib().hgb_nbody_sym_counters(nb._h, c))
c = list(c)
print(f"n={n} sym 1242: {ms:.1f} ms; of worker time: wait {100 * c[0] / c[5]:.1f} % open {100 * c[1] / c[5]:.1f} % close {100 * c[2] / c[5]:.1f} % "
      f"fast loop {100 * c[3] / c[5]:.1f} % edge groups {100 * c[4] / c[5]:.1f} %; chunks {c[6]}, fast-loop cycles per chunk {c[3] / c[6]:.0f}")
T = (n + 63) // 64
print(f"raw: wait {c[0]:.4e} open {c[1]:.4e} close {c[2]:.4e} fast {c[3]:.4e} edge {c[4]:.4e} work {c[5]:.4e} chunks {c[6]}; workers {T}; "
      f"edge-group cycles per worker {c[4] / T:.0f} (32 groups: {c[4] / T / 32:.0f} each); wait per chunk {c[0] / c[6]:.0f}; "
      f"per worker: total {c[5] / T:.4e}")
H.lib().hgb_finalize()
