#!/usr/bin/env python
"""Parity of the ONE-PROCESS-PER-GPU path (what bench.py / the driver's scaling run use) against the
oracle.  Launch under torchrun:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 \
      --master-port P tools/rank_mode_check.py

Every rank builds rank-mode handles (hgb_nbody_create_rank / hgb_conv_create_rank over the NCCL
communicator of hgb_comm_init_rank), runs the sharded step (ncclAllGather of positions; row bands
with ncclSend/ncclRecv halos, deep-halo cycles, boundary/interior overlap) and compares its result
with the oracle bit for bit.  Exit code != 0 on any mismatch.  tests/test_rank_mode_gpu.py wraps it.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    import hg_c_benchmarks_b200 as H
    from hg_c_benchmarks_b200 import synth
    from oracle import port

    rank, world = H.dist.init_rank_comm()
    checks = 0

    # ---- nbody: target slices + ncclAllGather ------------------------------------------------
    for n, iters in ((5000, 3), (2048, 2), (70001, 2), (1 << 17, 1)):
        pos = synth.plummer(n, seed=n)
        with H.nbody.NBody(n, ngpus=None, mode=H.NBODY_EXACT) as nb:
            nb.upload(pos)
            nb.step(iters)
            cur, prev = nb.download(want_prev=True)
        if n <= 5000:        # the whole run
            a, b = pos.copy(), np.zeros_like(pos)
            port.nbody_run(a, b, iters)
            w_cur, w_prev = (a, b) if iters % 2 == 0 else (b, a)
            assert np.array_equal(cur, w_cur) and np.array_equal(prev, w_prev), ("nbody run", n, rank)
        # the last step on a target sample that covers every rank's slice, slice seams included
        s = H.dist.nbody_slice(n, world)
        seams = [t for r in range(world) for t in (r * s - 1, r * s, r * s + 1) if 0 <= t < n]
        sample = np.unique(np.concatenate([np.random.default_rng(n).choice(n, 160, replace=False),
                                           np.array(seams + [0, n - 1])])).astype(np.int32)
        assert np.array_equal(cur[sample], port.nbody_accel_subset(prev, sample)), ("nbody sample", n, rank)
        checks += 1

    # ---- nbody: pair-symmetric wavefront, tile-rows dealt over the ranks, chain words through CUDA IPC mappings ----
    tri_kernels = set()
    # second generation (one warp per row) and third (a CTA of four warps per row): same buffers, same dealing
    for variant, n, iters, seg, rows, weight in ((2002, 5000, 3, 256, 2, 0), (2002, 2048, 5, 64, 1, 4), (2002, 70001, 2, 4096, 4, 0),
                                                 (2002, 1 << 17, 3, 0, 0, 0), (2002, 33, 2, 32, 1, 1), (2002, 1 << 15, 4, 1024, 8, 2),
                                                 (3002, 5000, 3, 256, 2, 0), (3002, 2048, 5, 96, 1, 4), (3002, 70001, 2, 4096, 4, 0),
                                                 (3002, 1 << 17, 3, 0, 0, 0), (3002, 33, 2, 32, 1, 1), (3002, 1 << 15, 4, 1056, 8, 2),
                                                 (6242, 5000, 3, 256, 2, 0), (6242, 2048, 5, 96, 1, 4), (6242, 70001, 2, 4096, 4, 0),
                                                 (6242, 1 << 17, 3, 0, 0, 0), (6242, 65, 2, 32, 1, 1), (6242, 1 << 15, 4, 1056, 8, 2)):
        pos = synth.plummer(n, seed=n + 1)
        with H.nbody.NBody(n, ngpus=None, mode=H.NBODY_EXACT) as nb:
            nb.symmetric(variant, 0).segment(seg, rows, weight)
            nb.upload(pos)
            nb.step(iters)
            cur, prev = nb.download(want_prev=True)
            tri_kernels.add(nb.describe()[0].split("<")[0])
        if n <= 5000:
            a, b = pos.copy(), np.zeros_like(pos)
            port.nbody_run(a, b, iters)
            w_cur, w_prev = (a, b) if iters % 2 == 0 else (b, a)
            assert np.array_equal(cur, w_cur) and np.array_equal(prev, w_prev), ("nbody tri run", n, rank)
        sample = np.unique(np.concatenate([np.random.default_rng(n).choice(n, min(n, 200), replace=False),
                                           np.array([0, 1, 31, 32, n // 2, n - 2, n - 1])])).astype(np.int32)
        assert np.array_equal(cur[sample], port.nbody_accel_subset(prev, sample)), ("nbody tri sample", n, rank)
        checks += 1
    # the default policy on a set large enough for every rank count: whatever kernel it picks, same bits
    n = 131072 * max(1, world)
    if n <= (1 << 20):
        pos = synth.uniform_cube(n, seed=5)
        with H.nbody.NBody(n, ngpus=None, mode=H.NBODY_EXACT) as nb:
            nb.upload(pos)
            nb.step(2)
            cur, prev = nb.download(want_prev=True)
            tri_kernels.add("default:" + nb.describe()[0].split(" ")[0])
        sample = np.random.default_rng(7).choice(n, 96, replace=False).astype(np.int32)
        assert np.array_equal(cur[sample], port.nbody_accel_subset(prev, sample)), ("nbody default policy", n, rank)
        checks += 1
    print(f"rank {rank}/{world}: pair-symmetric cases ran {sorted(tri_kernels)}", flush=True)

    # ---- convolution: row bands + halo exchange ------------------------------------------------
    for w, h, k, it, depth in ((300, 70, 3, 4, 0), (513, 260, 5, 3, 1), (640, 33, 7, 3, 0), (64, 5, 7, 2, 0),
                               (300, 1100, 3, 21, 0), (513, 1300, 7, 9, 3), (260, 2049, 5, 10, 0),
                               (2048, 4096, 3, 10, 0), (1024, 2048, 7, 10, 0)):
        p = k // 2
        kern = synth.kernel_dense(k)
        img = synth.add_padding(synth.image_float(w, h, seed=w + h), p)
        want_c, want_p = img.copy(), np.zeros_like(img)
        port.conv_run(want_c, want_p, w, h, p, kern, it)
        with H.convolution.Conv2D(w, h, p, ngpus=None) as cv:
            r0, rows = cv.band()
            cv.set_kernel(kern)
            cv.set_depth(depth)
            cv.upload(np.ascontiguousarray(img[r0:r0 + rows]))
            cv.step(it)
            cur, prev = cv.download(want_prev=True)
        assert np.array_equal(cur, want_c[r0:r0 + rows]), ("conv cur", w, h, k, it, rank)
        assert np.array_equal(prev, want_p[r0:r0 + rows]), ("conv prev", w, h, k, it, rank)
        checks += 1
    print(f"RANK_MODE_OK rank {rank}/{world}: {checks} cases bit-identical to the oracle", flush=True)
    import torch
    if world > 1:
        torch.distributed.barrier()
    H.lib().hgb_finalize()
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
