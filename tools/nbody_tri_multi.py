"""nbody_tri_kernel across the GPUs of one box (one process driving g GPUs, peer-mapped buffers): device
time per step beside the standard kernel with its allgather, results compared with the 1-GPU standard
kernel.  usage: python tools/nbody_tri_multi.py [n] [g ...]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
gs = [int(a) for a in sys.argv[2:]] or [g for g in (1, 2, 4, 8) if g <= torch.cuda.device_count()]
pos = synth.uniform_cube(n, seed=1)


def run(g, variant, seg=0, rows=0, weight=0, steps=2, deal=0):
    with H.nbody.NBody(n, ngpus=g, exchange=H._abi.EXCHANGE_P2P if variant else H._abi.EXCHANGE_NCCL) as nb:
        nb.symmetric(variant, 0)
        if variant >= 2000:
            nb.segment(seg, rows, weight).deal(deal)
        nb.upload(pos)
        nb.step(1)
        nb.upload(pos)
        ms = nb.step(steps) / steps
        out = nb.download()
        counters = None
        if variant >= 2010 and (variant // 10) % 10 == 1:
            c = (C.c_ulonglong * 7)()
            H._abi.check(H.lib().hgb_nbody_sym_counters(nb._h, c))
            counters = list(c)
        what = nb.describe()
    return ms, out, counters, what


ref = None
for g in gs:
    if g == 1 and len(gs) > 1:
        ms, out, _, what = run(1, 0)
        ref = out
        print(f"g=1 n={n} standard: {ms:9.3f} ms", flush=True)
        continue
    ms, out, _, what = run(g, 0)
    if ref is None:
        ref = out
    print(f"g={g} n={n} standard: {ms:9.3f} ms  {n * n / (ms * 1e-3):.4e} inter/s  same={np.array_equal(out, ref)}  [{what}]", flush=True)
    grid = [(2002, 16384, r, w) for r in (16, 32, 64) for w in (32, 64, 128, 256)] + [(2002, 8192, r, w) for r in (16, 32) for w in (64, 128)]
    if os.environ.get("TRI_GRID") == "small":
        grid = [(2002, 16384, 8, 32), (2002, 16384, 16, 32), (2002, 16384, 16, 64), (2002, 8192, 16, 32), (2003, 16384, 16, 32)]
    if os.environ.get("TRI_GRID") == "policy4":
        grid = [(2002, 16384, 16, 64, 16), (2002, 16384, 8, 64, 8), (3002, 0, 0, 0, 0), (3002, 16384, 8, 32, 16), (6242, 0, 0, 0, 0)]
    if os.environ.get("TRI_GRID") == "seg":
        grid = [(2002, 16384, 16, 64, 16), (6242, 0, 0, 0, 0), (6242, 16384, 2, 128, 8), (6242, 16384, 8, 32, 8), (6242, 16384, 4, 64, 16),
                (6242, 32768, 4, 64, 8), (6242, 16384, 4, 128, 4), (3002, 0, 0, 0, 0)]
    if os.environ.get("TRI_GRID") == "crow":
        grid = [(3002, 0, 0, 0, 0), (3002, 16384, 8, 32, 16), (3002, 16384, 4, 64, 8), (3002, 16384, 8, 16, 8), (3002, 16384, 8, 64, 8),
                (3002, 16384, 4, 32, 4), (3002, 8192, 8, 64, 8), (3002, 32768, 8, 32, 8), (3002, 16384, 4, 32, 8)]
    for spec in grid:
        variant, seg, rows, weight = spec[:4]
        deal = spec[4] if len(spec) > 4 else 0
        ms, out, _, what = run(g, variant, seg, rows, weight, deal=deal)
        print(f"g={g} n={n} {'crow' if variant >= 3000 else 'tri'} {variant} seg={seg:6d} rows={rows:3d} weight={weight:3d} deal={deal:3d}: {ms:9.3f} ms  "
              f"{n * n / (ms * 1e-3):.4e} inter/s  same={np.array_equal(out, ref)}  [{what[0]}]", flush=True)
    crow = os.environ.get("TRI_GRID") == "crow"
    if os.environ.get("TRI_GRID") in ("seg", "policy4"):
        continue
    ms, out, c, _ = run(g, 3012 if crow else 2012, 16384, 8 if crow else 16, 32 if crow else 64, steps=1, deal=8 if crow else 16)
    tot = max(1, c[2])
    print(f"g={g} n={n} tri 2012 (profiling, shard 0): {ms:9.3f} ms; chain-word waits {100.0 * c[0] / tot:.1f} %, "
          f"row-accumulator waits {100.0 * c[1] / tot:.1f} %, items {c[4]}, cycles per item {tot / max(1, c[4]):.0f}", flush=True)
H.lib().hgb_finalize()
