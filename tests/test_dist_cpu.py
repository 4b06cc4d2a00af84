"""Multi-rank host logic on CPU: world_size-2 (and 3) gloo groups run the SAME shard plans
the library uses (hg_c_benchmarks_b200.dist), with the oracle standing in for the kernels
-- here only, as the checker -- and the result must equal the single-rank oracle bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hg_c_benchmarks_b200 import dist as plan
from hg_c_benchmarks_b200 import synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _nbody_rank(rank, world, port_no, n, iters, q):
    from oracle import port
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port_no}", rank=rank, world_size=world)
    pos = synth.plummer(n, seed=11)
    sl = plan.nbody_slice(n, world)
    npad = sl * world
    # SoA ping-pong buffers with the library's padded layout
    cur = np.zeros((3, npad))
    cur[:, :n] = pos.T
    for _ in range(iters):
        t0, t1 = plan.nbody_targets(n, world, rank)
        aos = np.ascontiguousarray(cur[:, :n].T)
        out = np.full((n, 3), np.nan)
        port.nbody_step_range(aos, out, t0, t1, threads=1)
        nxt = np.full((3, npad), np.nan)
        nxt[:, t0:t1] = out[t0:t1].T
        for c in range(3):   # in-place style all-gather of equal padded slices, one per component
            send = torch.from_numpy(nxt[c, rank * sl:(rank + 1) * sl].copy())
            recv = [torch.empty(sl, dtype=torch.float64) for _ in range(world)]
            dist.all_gather(recv, send)
            nxt[c] = torch.cat(recv).numpy()
        cur = nxt
    if rank == 0:
        q.put(np.ascontiguousarray(cur[:, :n].T))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n,iters", [(2, 300, 3), (2, 2049, 2), (3, 1000, 2)])
def test_nbody_target_sharding_gloo(world, n, iters):
    from oracle import port
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port_no = _free_port()
    procs = [ctx.Process(target=_nbody_rank, args=(r, world, port_no, n, iters, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    a, b = synth.plummer(n, seed=11), np.zeros((n, 3))
    port.nbody_run(a, b, iters)
    want = a if iters % 2 == 0 else b
    assert np.array_equal(got, want)


def _conv_rank(rank, world, port_no, w, h, p, k, iters, q):
    from oracle import port
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port_no}", rank=rank, world_size=world)
    kern = synth.kernel_dense(k)
    full = synth.add_padding(synth.image_float(w, h, seed=5), p)
    r0, rows = plan.conv_band(h, p, world, rank)
    # rows outside band+halo are poisoned: a wrong halo plan shows up as NaN
    cur = np.full_like(full, np.nan)
    lo, hi = max(0, r0 - p), min(h, r0 + rows + p)
    if rows:
        cur[lo:hi] = full[lo:hi]
    for _ in range(iters):
        nxt = np.full_like(full, np.nan)
        if rows:
            port.conv_step_rows(cur, nxt, w, h, p, kern, r0, r0 + rows, threads=1)
        reqs = []
        bufs = []
        for peer, (s0, s1), (v0, v1) in plan.conv_halo_plan(h, p, world, rank):
            reqs.append(dist.isend(torch.from_numpy(nxt[s0:s1].copy()), peer))
            t = torch.empty((v1 - v0, w + 2 * p), dtype=torch.float64)
            bufs.append((v0, v1, t))
            reqs.append(dist.irecv(t, peer))
        for r in reqs:
            r.wait()
        for v0, v1, t in bufs:
            nxt[v0:v1] = t.numpy()
        cur = nxt
    q.put((rank, r0, rows, cur[r0:r0 + rows].copy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h,k,iters", [(2, 40, 31, 3, 4), (2, 33, 20, 7, 3), (3, 25, 8, 5, 3), (2, 16, 2, 7, 2)])
def test_conv_row_bands_with_halo_exchange_gloo(world, w, h, k, iters):
    from oracle import port
    p = k // 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port_no = _free_port()
    procs = [ctx.Process(target=_conv_rank, args=(r, world, port_no, w, h, p, k, iters, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    parts = [q.get(timeout=300) for _ in range(world)]
    for pr in procs:
        pr.join(timeout=120)
        assert pr.exitcode == 0
    want = synth.add_padding(synth.image_float(w, h, seed=5), p)
    prev = np.zeros_like(want)
    port.conv_run(want, prev, w, h, p, synth.kernel_dense(k), iters)
    got = np.full_like(want, np.nan)
    for _, r0, rows, band in parts:
        got[r0:r0 + rows] = band
    assert np.array_equal(got, want)


def _conv_deep_rank(rank, world, port_no, w, h, p, k, iters, depth, q):
    """the communication-avoiding schedule of csrc/hgb_conv.cu hgb_conv_step on CPU: depth*p halo rows
    travel once per cycle of `depth` iterations, the shrinking fringe is recomputed in between"""
    from oracle import port
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port_no}", rank=rank, world_size=world)
    kern = synth.kernel_dense(k)
    full = synth.add_padding(synth.image_float(w, h, seed=5), p)
    r0, rows = plan.conv_band(h, p, world, rank)
    d_eff = plan.conv_eff_depth(h, p, world, depth)
    xr = d_eff * p
    cur = np.full_like(full, np.nan)          # rows outside band + halo are poison
    if rows:
        cur[max(0, r0 - xr):min(h, r0 + rows + xr)] = full[max(0, r0 - xr):min(h, r0 + rows + xr)]
    exchanges = 0
    it = 0
    while it < iters:
        d = min(d_eff, iters - it)
        for t in range(d):
            nxt = np.full_like(full, np.nan)
            lo, hi = plan.conv_cycle_rows(h, p, world, rank, d, t)
            if rows:
                port.conv_step_rows(cur, nxt, w, h, p, kern, lo, hi, threads=1)
            cur = nxt
            it += 1
        reqs, bufs = [], []
        for peer, (s0, s1), (v0, v1) in plan.conv_deep_halo_plan(h, p, world, rank, d_eff):
            reqs.append(dist.isend(torch.from_numpy(cur[s0:s1].copy()), peer))
            tbuf = torch.empty((v1 - v0, w + 2 * p), dtype=torch.float64)
            bufs.append((v0, v1, tbuf))
            reqs.append(dist.irecv(tbuf, peer))
        for r in reqs:
            r.wait()
        for v0, v1, tbuf in bufs:
            cur[v0:v1] = tbuf.numpy()
        exchanges += 1
    q.put((rank, r0, rows, cur[r0:r0 + rows].copy(), exchanges, d_eff))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h,k,iters,depth", [(2, 40, 31, 3, 7, 3), (2, 33, 40, 7, 5, 2), (3, 25, 30, 5, 6, 0),
                                                     (2, 24, 9, 3, 5, 8), (2, 16, 6, 7, 3, 4)])
def test_conv_deep_halo_schedule_gloo(world, w, h, k, iters, depth):
    from oracle import port
    p = k // 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port_no = _free_port()
    procs = [ctx.Process(target=_conv_deep_rank, args=(r, world, port_no, w, h, p, k, iters, depth, q))
             for r in range(world)]
    for pr in procs:
        pr.start()
    parts = [q.get(timeout=300) for _ in range(world)]
    for pr in procs:
        pr.join(timeout=120)
        assert pr.exitcode == 0
    want = synth.add_padding(synth.image_float(w, h, seed=5), p)
    prev = np.zeros_like(want)
    port.conv_run(want, prev, w, h, p, synth.kernel_dense(k), iters)
    got = np.full_like(want, np.nan)
    for _, r0, rows, band, exchanges, d_eff in parts:
        got[r0:r0 + rows] = band
        assert exchanges == -(-iters // d_eff)            # one exchange per cycle, not per iteration
        assert 1 <= d_eff <= max(1, (h // plan.conv_active_ranks(h, p, world)) // p)
    assert np.array_equal(got, want)


def test_conv_eff_depth_clamps():
    assert plan.conv_eff_depth(16384, 1, 8) == 8                    # automatic
    assert plan.conv_eff_depth(16384, 1, 8, fused3x3=True) == 16    # 3x3 with two iterations per pass
    assert plan.conv_eff_depth(16384, 3, 8) == 8
    assert plan.conv_eff_depth(16384, 3, 8, depth=20) == 10         # halo capacity: 32 rows / p
    assert plan.conv_eff_depth(40, 1, 8, depth=16) == 5             # shortest band: 5 rows
    assert plan.conv_eff_depth(16384, 1, 1, depth=16) == 1          # one band: nothing to exchange
    assert plan.conv_eff_depth(6, 3, 2, depth=4) == 1


def test_plans_cover_everything_once():
    for n in (1, 3, 1000, 1024, 1025, 65536, 1 << 20):
        for world in (1, 2, 3, 4, 8):
            sl = plan.nbody_slice(n, world)
            assert sl % plan.NBODY_ALIGN == 0 and sl * world >= n
            covered = []
            for r in range(world):
                t0, t1 = plan.nbody_targets(n, world, r)
                covered += list(range(t0, t1)) if n <= 2000 else []
                assert 0 <= t0 <= t1 <= n
            if n <= 2000:
                assert covered == list(range(n))
            assert sum(plan.nbody_targets(n, world, r)[1] - plan.nbody_targets(n, world, r)[0]
                       for r in range(world)) == n
    for h in (1, 2, 7, 512, 16384):
        for p in (0, 1, 2, 3):
            for world in (1, 2, 4, 8):
                act = plan.conv_active_ranks(h, p, world)
                rows = [plan.conv_band(h, p, world, r) for r in range(world)]
                assert sum(r[1] for r in rows) == h
                nxt = 0
                for r0, cnt in rows[:act]:
                    assert r0 == nxt and (act == 1 or cnt >= max(p, 1))
                    nxt += cnt
                for r in range(world):
                    for peer, (s0, s1), (v0, v1) in plan.conv_halo_plan(h, p, world, r):
                        pr0, prows = plan.conv_band(h, p, world, peer)
                        assert pr0 <= v0 and v1 <= pr0 + prows      # what I receive is the peer's own rows
                        assert s1 - s0 == p and v1 - v0 == p


def test_env_rank_defaults(monkeypatch):
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    assert plan.env_rank() == (0, 0, 1)
    monkeypatch.setenv("RANK", "3")
    monkeypatch.setenv("LOCAL_RANK", "1")
    monkeypatch.setenv("WORLD_SIZE", "4")
    assert plan.env_rank() == (3, 1, 4)
