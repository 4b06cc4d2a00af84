"""Where the warps of the pair-symmetric nbody kernel spend their time: clock64 spans of the profiling
instantiations (variant + 1000), summed over all warps.  usage: python tools/nbody_sym_prof.py [n ...]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

NAMES = ("flag wait", "chunk loads", "chunk release", "fast loop", "predicated steps", "workers total", "chunks")
for n in [int(a) for a in sys.argv[1:]] or [65536, 1 << 20]:
    pos = synth.uniform_cube(n, seed=1)
    for v in (1226, 1242):
        with H.nbody.NBody(n) as nb:
            nb.symmetric(v, 0).upload(pos)
            nb.step(1)
            nb.upload(pos)
            ms = nb.step(1)
            cnt = (C.c_ulonglong * 7)()
            H._abi.check(H.lib().hgb_nbody_sym_counters(nb._h, cnt))
        warps = 148 * (v % 10) * 4
        print(f"n={n} variant={v}: {ms:.3f} ms, {warps} warps", flush=True)
        chunks = max(1, cnt[6])
        for name, c in zip(NAMES, cnt):
            if name == "chunks":
                print(f"    {name:18s} {c}")
            else:
                print(f"    {name:18s} {c / 1e6:12.1f} Mcycles  {100.0 * c / max(1, cnt[5]):6.2f} % of worker time  "
                      f"{c / chunks:10.1f} cycles per chunk")
H.lib().hgb_finalize()
