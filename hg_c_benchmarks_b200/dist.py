"""Multi-GPU host logic: shard plans and the one-process-per-GPU communicator.

The plans here are the SAME arithmetic libhgb200.so uses (csrc/hgb_nbody.cu
`shard_size`, csrc/hgb_conv.cu `band_of`); tests/test_dist_cpu.py checks them
on CPU with a world_size-2 gloo group.  torch.distributed is plumbing only
(rendezvous, broadcast of the NCCL id, max-over-ranks of timings).
"""
import os

NBODY_ALIGN = 1024


def nbody_slice(n, world, align=NBODY_ALIGN):
    """padded per-rank target slice; rank r owns targets [r*slice, min(n, (r+1)*slice))"""
    per = (n + world - 1) // world
    return (per + align - 1) // align * align


def nbody_targets(n, world, rank):
    s = nbody_slice(n, world)
    t0 = min(n, rank * s)
    return t0, min(n, t0 + s)


def conv_active_ranks(h, p, world):
    """bands must be >= p rows tall so halos only come from direct neighbours"""
    need = p if p > 0 else 1
    return max(1, min(world, h // need))


def conv_band(h, p, world, rank):
    """(row0, rows) of rank's band; ranks >= active own nothing"""
    act = conv_active_ranks(h, p, world)
    if rank >= act:
        return h, 0
    base, rem = divmod(h, act)
    return rank * base + min(rank, rem), base + (1 if rank < rem else 0)


def conv_halo_plan(h, p, world, rank):
    """list of (peer, send_rows(lo,hi) in global rows, recv_rows(lo,hi)) for one exchange"""
    act = conv_active_ranks(h, p, world)
    r0, rows = conv_band(h, p, world, rank)
    plan = []
    if rows == 0 or p == 0:
        return plan
    if rank > 0:
        plan.append((rank - 1, (r0, r0 + p), (r0 - p, r0)))
    if rank < act - 1:
        plan.append((rank + 1, (r0 + rows - p, r0 + rows), (r0 + rows, r0 + rows + p)))
    return plan


CONV_HALO_ROWS = 32   # csrc/hgb_conv.cu kBnd: halo capacity per side of a band


def conv_eff_depth(h, p, world, depth=0, fused3x3=False):
    """iterations per halo exchange (csrc/hgb_conv.cu eff_depth): the request (0 = automatic: 8, or
    16 where 3x3 passes run two iterations each) clamped by the halo capacity and the shortest band"""
    act = conv_active_ranks(h, p, world)
    if act <= 1 or p <= 0:
        return 1
    halo = CONV_HALO_ROWS if p < CONV_HALO_ROWS else p
    d = depth if depth > 0 else (16 if (p == 1 and fused3x3) else 8)
    d = min(d, halo // p, (h // act) // p)
    return max(1, d)


def conv_cycle_rows(h, p, world, rank, d, t):
    """global rows [lo, hi) that `rank` computes in iteration t of a cycle of d iterations: its
    band plus the shrinking fringe of (d-1-t)*p rows per side (communication-avoiding halo)"""
    r0, rows = conv_band(h, p, world, rank)
    if rows == 0:
        return h, h
    ext = (d - 1 - t) * p
    return max(0, r0 - ext), min(h, r0 + rows + ext)


def conv_deep_halo_plan(h, p, world, rank, d):
    """one exchange of d*p rows per side after a cycle: (peer, send(lo,hi), recv(lo,hi)), global rows"""
    act = conv_active_ranks(h, p, world)
    r0, rows = conv_band(h, p, world, rank)
    xr = d * p
    plan = []
    if rows == 0 or p == 0:
        return plan
    if rank > 0:
        plan.append((rank - 1, (r0, r0 + xr), (r0 - xr, r0)))
    if rank < act - 1:
        plan.append((rank + 1, (r0 + rows - xr, r0 + rows), (r0 + rows, r0 + rows + xr)))
    return plan


def env_rank():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def init_rank_comm(device=None):
    """One process per GPU (torchrun).  Selects the GPU, and when WORLD_SIZE > 1 brings up
    torch.distributed (NCCL) and the library's own NCCL communicator from a broadcast id."""
    import ctypes as C

    import torch

    from ._abi import UNIQUE_ID_BYTES, check, lib

    rank, local_rank, world = env_rank()
    dev = local_rank if device is None else device
    torch.cuda.set_device(dev)
    torch.cuda.init()
    torch.zeros(1, device=f"cuda:{dev}")      # make the primary context current
    buf = (C.c_ubyte * UNIQUE_ID_BYTES)()
    if world > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device(f"cuda:{dev}"))
        idt = torch.zeros(UNIQUE_ID_BYTES, dtype=torch.uint8)
        if rank == 0:
            check(lib().hgb_comm_unique_id(buf, UNIQUE_ID_BYTES))
            idt = torch.tensor(list(bytes(buf)), dtype=torch.uint8)
        idt = idt.cuda()
        dist.broadcast(idt, src=0)
        raw = bytes(idt.cpu().tolist())
        buf = (C.c_ubyte * UNIQUE_ID_BYTES).from_buffer_copy(raw)
    check(lib().hgb_comm_init_rank(buf, UNIQUE_ID_BYTES, rank, world))
    return rank, world
