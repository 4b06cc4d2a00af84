#!/usr/bin/env python
"""bench.py -- headline benchmark of the hg-c-benchmarks hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One "step" = one all-pairs FP64 nbody step over N = 2^20 bodies (BASELINE.json configs[3]),
EXACT mode (bit-identical to the reference).  Rank 0 prints ONE JSON line.  The same line carries
  * "parity_check": after the timed region every rank compares its results with the oracle (a target
    sample of the last nbody step; row bands of the convolution results incl. the band seams) --
    a mismatch makes the run exit non-zero;
  * "nbody_64k_plummer_100" (configs[2]), "nbody_1m_plummer" (SURVEY 8d item 4);
  * "conv": 16384^2 images, 3x3 / 7x7 dense (configs[4]), the shipped 5x5 kernel.txt, an integer image;
  * "cpu_baseline" (N=1): the unmodified reference on the host cores.
`--impl reference` times the unmodified reference (oracle/_ref, else the oracle port) on the host
cores for the same metric.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

OUT = sys.stdout
N_BODIES = 1 << 20
CONV_N = 16384
CONV_ITERS = 10          # the config's iteration count (used by the end-to-end leg and its parity check)
CONV_TIMED = 100         # device-resident timing runs 10 x 10 iterations back to back: at 8 GPUs ten
                         # iterations last ~1 ms, comparable to the host skew between ranks after a barrier
FLOP_PER_INTERACTION = 19          # 3 sub, 3 mul+3 add (+EPS), 2 mul, sqrt, div, 3 mul+3 add
FP64_INSTR_PER_INTERACTION = 29    # exact mode: 17 plain + 8 (IEEE sqrt) + 4 (IEEE reciprocal) FP64-pipe ops (tools/sass_loop.py)
BYTES_PER_PIXEL = 16               # 8 B read + 8 B written per inner pixel per iteration
PARITY_TARGETS = 256
PARITY_BAND = 64


def workload(n):
    """config.workload, shared by both arms (the reference arm runs a bounded sample of it)"""
    return (f"nbody N={n} uniform cube seed 1, all-pairs FP64, EXACT mode (bit-identical to the reference), "
            "BASELINE configs[3]")


def env_rank():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace(".", "").isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for name, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        # median over the samples taken under load (upper half of the sorted clocks is robust to idle gaps)
        med = sm[len(sm) // 2] if sm else None
        pw = [float(r[3]) for r in self.rows if len(r) > 3 and r[3].replace(".", "").isdigit()]
        return {"sm_mhz": med, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw) if pw else None}


def ncu_traffic(kernel_prefix):
    """DRAM bytes per launch of a kernel from the committed ncu --set full capture
    (profiles/ncu_traffic.json, written by tools/ncu_summary.py --traffic); None if absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            table = json.load(f)
    except (OSError, ValueError):
        return None
    for name, row in table.items():
        if name.startswith(kernel_prefix):
            return row["dram_bytes_per_launch"]
    return None


def pcie_probe(device):
    """pinned-memory copy rates of this box, both directions AT ONCE (what a streamed run can reach at best):
    ms per GiB for H2D and D2H, CUDA events on two torch streams"""
    import torch
    n = 1 << 27                                               # 1 GiB of doubles
    a_h, b_h = torch.empty(n, dtype=torch.float64).pin_memory(), torch.empty(n, dtype=torch.float64).pin_memory()
    a_d, b_d = torch.empty(n, dtype=torch.float64, device=device), torch.zeros(n, dtype=torch.float64, device=device)
    s1, s2 = torch.cuda.Stream(device), torch.cuda.Stream(device)
    res = {}
    for rep in range(2):
        torch.cuda.synchronize(device)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        with torch.cuda.stream(s1):
            e[0].record(); a_d.copy_(a_h, non_blocking=True); e[1].record()
        with torch.cuda.stream(s2):
            e[2].record(); b_h.copy_(b_d, non_blocking=True); e[3].record()
        torch.cuda.synchronize(device)
        res = {"h2d_ms_per_gib": e[0].elapsed_time(e[1]), "d2h_ms_per_gib": e[2].elapsed_time(e[3])}
    res["h2d_GBps"] = 2 ** 30 / (res["h2d_ms_per_gib"] * 1e-3) / 1e9
    res["d2h_GBps"] = 2 ** 30 / (res["d2h_ms_per_gib"] * 1e-3) / 1e9
    res["what"] = "1 GiB pinned host <-> device, both directions concurrently"
    del a_h, b_h, a_d, b_d
    return res


def pinned(shape):
    """numpy array backed by pinned host memory (the e2e legs copy from/to pinned buffers)"""
    import torch
    t = torch.empty(shape, dtype=torch.float64, pin_memory=True)
    return t.numpy(), t


# ---------------------------------------------------------------------------------------------
# CPU side: the unmodified reference (oracle/_ref) where it was built, else the oracle port
def reference_builds(step, n_cal=8192):
    """interactions/s of every build of the reference that runs on this host (one warm + one timed
    step over n_cal bodies); {} when only the port exists"""
    from hg_c_benchmarks_b200 import synth
    from oracle import ref
    rates = {}
    if not ref.available():
        return rates
    cal = synth.uniform_cube(n_cal, seed=1)
    buf = np.zeros_like(cal)
    for variant in ref.VARIANTS:
        if not ref.available(variant):
            continue
        ref.use(variant)
        step(cal, buf)
        t0 = time.perf_counter()
        step(cal, buf)
        rates[variant] = float(n_cal) ** 2 / (time.perf_counter() - t0)
    ref.use("")
    return rates


def run_reference(args):
    """the reference's own CPU implementation on the host cores, bounded sample"""
    rank, _, world = env_rank()
    if rank != 0:
        return
    from hg_c_benchmarks_b200 import synth
    from oracle import port, ref
    cores = port.ncpu()
    use_ref = ref.available()
    kind = "reference" if use_ref else "port"

    def step(pos, buf):
        if use_ref:
            ref.nbody_run(pos, buf, 1, cores)
        else:
            port.nbody_run(pos, buf, 1, cores)

    # every build of the unmodified reference that runs here is calibrated and the fastest is timed:
    # shipped flags (x86-64-v3), -O3 + profile feedback (what pgo.sh intended, oracle/Makefile ref-pgo),
    # and the AVX-512 build (the reference's -march=native on such a host)
    builds = reference_builds(step)
    if builds:
        best = max(builds, key=builds.get)
        ref.use(best)
        rate, build_name = builds[best], ref.VARIANTS[best]
    else:
        cal = synth.uniform_cube(8192, seed=1)
        buf = np.zeros_like(cal)
        step(cal, buf)
        t0 = time.perf_counter()
        step(cal, buf)
        rate, build_name = 8192.0 ** 2 / (time.perf_counter() - t0), "oracle port (oracle/hg_oracle.c), -O2"
    budget_s = 150.0 / max(1, args.steps + args.warmup)
    n = int(min(N_BODIES, max(8192, (rate * min(6.0, budget_s)) ** 0.5))) // 1024 * 1024
    pos = synth.uniform_cube(N_BODIES, seed=1)[:n].copy()     # first n bodies of the same workload
    buf = np.zeros_like(pos)
    for _ in range(args.warmup):
        step(pos.copy(), buf)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(pos.copy(), buf)
    dt = (time.perf_counter() - t0) / args.steps
    value = float(n) * n / dt
    sample = (f"first {n} of the {N_BODIES} bodies, 1 step = {n}^2 interactions per step, "
              f"{args.steps} steps, OpenMP over {cores} host threads; build: {build_name}")
    caveat = (f" [CPU arm: each step is a bounded sample, the first {n} bodies = {n}^2 interactions; the rate "
              f"(interactions/s) is size-independent for this O(N^2) loop, a full N={N_BODIES} step would take "
              f"{float(N_BODIES) ** 2 / value:.0f} s on these {cores} threads]")
    line = {"impl": "reference", "metric": "nbody_fp64_interactions_per_s", "value": value,
            "unit": "interactions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload(N_BODIES) + caveat,
                       "sample": "CPU arm: a bounded sample of the workload per step, see cpu_baseline.sample"},
            "cpu_baseline": {"value": value, "unit": "interactions/s", "cores": cores, "kind": kind, "sample": sample,
                             "builds_calibrated": {ref.VARIANTS[k]: v for k, v in builds.items()}},
            "e2e": {"value": value, "unit": "interactions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=OUT, flush=True)


def cpu_baseline():
    """bounded CPU samples on rank 0 at N=1: the unmodified reference on all
    host cores and on one (-p 1); nbody; convolution 3x3 and 7x7 at 4096^2 and one 16384^2 iteration"""
    from hg_c_benchmarks_b200 import synth
    from oracle import port, ref
    cores = port.ncpu()
    use_ref = ref.available()

    def nb_run(a, b, procs):
        if use_ref:
            ref.nbody_run(a, b, 1, procs)
        else:
            port.nbody_run(a, b, 1, procs)

    def cv_run(img, prev, w, h, p, k, it, procs):
        if use_ref:
            ref.conv_run(img, prev, w, h, p, k, it, procs)
        else:
            port.conv_run(img, prev, w, h, p, k, it, procs)

    builds = reference_builds(lambda a, b: nb_run(a, b, cores))
    rate = builds.get("", None)
    if rate is None:
        cal = synth.uniform_cube(8192, seed=1)
        buf = np.zeros_like(cal)
        nb_run(cal, buf, cores)
        t0 = time.perf_counter()
        nb_run(cal, buf, cores)
        rate = 8192.0 ** 2 / (time.perf_counter() - t0)
    n = int(min(262144, max(16384, (rate * 10.0) ** 0.5))) // 1024 * 1024
    pos = synth.uniform_cube(N_BODIES, seed=1)[:n].copy()
    buf = np.zeros_like(pos)
    t0 = time.perf_counter()
    nb_run(pos, buf, cores)
    dt = time.perf_counter() - t0
    out = {"value": float(n) * n / dt, "unit": "interactions/s", "cores": cores,
           "kind": "reference" if use_ref else "port",
           "sample": f"first {n} of the {N_BODIES} bodies, one step ({n}^2 interactions, {dt:.1f} s), "
                     f"reference run() with number_of_processes={cores}",
           "build": "shipped effective flags -std=c11 -fcommon -O2 -fopenmp (CMakeLists.txt:40,45); -march=x86-64-v3 "
                    "instead of the reference's -march=native so that the binary runs on any host (the AVX-512 "
                    "and profile-feedback builds are calibrated under `builds`)"}
    if builds:
        out["builds"] = {ref.VARIANTS[k]: {"value": v, "unit": "interactions/s", "sample": "8192 bodies, one step"}
                         for k, v in builds.items()}
    # -p 1: one host thread (the reference's default)
    m = 16384
    q, qb = pos[:m].copy(), np.zeros((m, 3))
    t0 = time.perf_counter()
    nb_run(q, qb, 1)
    d1 = time.perf_counter() - t0
    out["p1"] = {"value": float(m) * m / d1, "unit": "interactions/s", "cores": 1,
                 "sample": f"first {m} bodies, one step ({d1:.1f} s), number_of_processes=1"}
    # convolution: 4096^2, dense 3x3 and 7x7, 2 iterations, all cores; 3x3 also on one thread
    w = 4096
    dense = synth.image_float(w, w, seed=3)
    conv = {}
    for k in (3, 7):
        p = k // 2
        img = synth.add_padding(dense, p)
        prev = np.zeros_like(img)
        kern = synth.kernel_dense(k)
        t0 = time.perf_counter()
        cv_run(img, prev, w, w, p, kern, 2, cores)
        dtc = time.perf_counter() - t0
        conv[f"{k}x{k}"] = {"value": w * w * 2 * BYTES_PER_PIXEL / dtc / 1e9, "unit": "GB/s", "cores": cores,
                            "sample": f"{w}x{w} dense {k}x{k}, 2 iterations, run_on_padded_image with {cores} threads"}
    img = synth.add_padding(dense[:1024], 1)
    prev = np.zeros_like(img)
    t0 = time.perf_counter()
    cv_run(img, prev, w, 1024, 1, synth.kernel_dense(3), 2, 1)
    dtc = time.perf_counter() - t0
    conv["3x3_p1"] = {"value": w * 1024 * 2 * BYTES_PER_PIXEL / dtc / 1e9, "unit": "GB/s", "cores": 1,
                      "sample": f"{w}x1024 dense 3x3, 2 iterations, one thread"}
    del dense, img, prev
    # one iteration at the benchmark size (2 x 2 GiB padded images on the host)
    try:
        big = synth.add_padding(synth.image_float(CONV_N, CONV_N, seed=3), 1)
        prev = np.zeros_like(big)
        t0 = time.perf_counter()
        cv_run(big, prev, CONV_N, CONV_N, 1, synth.kernel_dense(3), 1, cores)
        dtc = time.perf_counter() - t0
        conv["3x3_16k"] = {"value": CONV_N * CONV_N * BYTES_PER_PIXEL / dtc / 1e9, "unit": "GB/s", "cores": cores,
                           "sample": f"{CONV_N}x{CONV_N} dense 3x3, ONE iteration ({dtc:.2f} s), {cores} threads"}
        del big, prev
    except MemoryError:
        conv["3x3_16k"] = None
    out["conv"] = conv["3x3"]
    out["conv_all"] = conv
    return out


# ---------------------------------------------------------------------------------------------
# parity: the oracle as the checker, outside every timed region
def nbody_parity(nb, n, world, rank, seed):
    """compare a target sample of the LAST step (state k from state k-1, both still on the device)
    with the oracle's accel(); returns (ok, targets checked)"""
    from hg_c_benchmarks_b200 import dist as hdist
    from oracle import port
    cur, prev = nb.download(want_prev=True)
    s = hdist.nbody_slice(n, world)
    seams = [t for r in range(world) for t in (r * s - 1, r * s) if 0 <= t < n]
    rng = np.random.default_rng(seed + 1000 * rank)
    sample = np.unique(np.concatenate([rng.choice(n, min(n, PARITY_TARGETS), replace=False),
                                       np.array(seams + [0, n - 1])])).astype(np.int32)
    want = port.nbody_accel_subset(prev, sample)
    return bool(np.array_equal(cur[sample], want)), int(sample.size)


def conv_parity(got, dense, r0, rows, cn, p, kern, iters, tol=0.0):
    """rows of this rank's band after `iters` iterations against the oracle run on a sub-image with an
    iters*p-row margin (pollution from the sub-image's artificial edges travels p rows per iteration;
    true image edges are the sub-image's own edges and clamp like the reference's).  Bands: the first
    PARITY_BAND rows of the band, its last ones (both touch the neighbours' halos = the band seams) and
    one in the middle.  tol == 0: bit-identical required (EXACT mode); tol > 0: max-norm relative error
    (FAST mode, north_star's 1e-12).  Returns (ok, rows checked, worst relative error)"""
    from hg_c_benchmarks_b200 import synth
    from oracle import port
    if rows <= 0:
        return True, 0, 0.0
    m = iters * p
    nb = min(PARITY_BAND, rows)
    starts = sorted({r0, r0 + max(0, rows // 2 - nb // 2), r0 + rows - nb})
    ok, checked, worst = True, 0, 0.0
    for a in starts:
        lo, hi = max(0, a - m), min(cn, a + nb + m)
        sub = synth.add_padding(dense[lo:hi], p)
        tmp = np.zeros_like(sub)
        port.conv_run(sub, tmp, cn, hi - lo, p, kern, iters)
        mine, want = got[a - r0:a - r0 + nb], sub[a - lo:a - lo + nb]
        err = float(np.abs(mine - want).max() / max(np.abs(want).max(), 1e-300))
        worst = max(worst, err)
        ok = ok and (bool(np.array_equal(mine, want)) if tol == 0.0 else err <= tol)
        checked += nb
    return ok, checked, worst


def main():
    # stdout carries exactly ONE JSON line: libraries that print to fd 1 (NCCL's version banner)
    # are sent to stderr, the JSON goes to a private duplicate of the original stdout
    global OUT
    sys.stdout.flush()
    OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bodies", type=int, default=N_BODIES, help=argparse.SUPPRESS)
    ap.add_argument("--conv-size", type=int, default=CONV_N, help=argparse.SUPPRESS)
    ap.add_argument("--skip-conv", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--skip-cpu", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--skip-extra", action="store_true", help=argparse.SUPPRESS)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference(args)

    import ctypes as C

    import torch

    import hg_c_benchmarks_b200 as H
    from hg_c_benchmarks_b200 import synth

    rank, local_rank, world = env_rank()
    L = H.lib()
    H.dist.init_rank_comm()
    dist = torch.distributed if world > 1 else None

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def all_ok(ok):
        if not dist:
            return bool(ok)
        t = torch.tensor([1.0 if ok else 0.0], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item() > 0.5)

    def total(x):
        if not dist:
            return x
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return int(t.item())

    peaks = measured_peaks()
    sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
    # FP64 issue-rate probe: ~1 s of dependent-free DFMAs after a warm-up, clocks sampled meanwhile
    probe_sampler = ClockSampler(local_rank).start() if rank == 0 else None
    tf, mhz, gbs = C.c_double(), C.c_double(), C.c_double()
    H._abi.check(L.hgb_measure_fp64_peak(C.byref(tf), C.byref(mhz)))
    probe_clocks = probe_sampler.stop() if probe_sampler else None
    H._abi.check(L.hgb_measure_hbm_copy(C.byref(gbs)))

    n = args.bodies
    pos_np, pos_t = pinned((n, 3))
    pos_np[...] = synth.uniform_cube(n, seed=1)
    out_np, out_t = pinned((n, 3))
    parity = {}

    # ---------------- nbody: device-resident steps -------------------------------------------
    nb = H.nbody.NBody(n, ngpus=None, mode=H.NBODY_EXACT)
    nb.upload(pos_np)
    for _ in range(args.warmup):
        nb.step(1)
    nb.step(0)                                   # the first call connects the ranks' buffers (and may fall back): ask afterwards
    sym = int(L.hgb_nbody_symmetric(nb._h))      # pair-symmetric variant in use (0: standard kernel)
    fp64_instr = float(L.hgb_nbody_fp64_instr(nb._h))
    nb_kernel, nb_parallelism = nb.describe()
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    t_wall0 = time.perf_counter()
    step_ms = []
    launches_timed = 0
    for _ in range(args.steps):
        H._abi.check(L.hgb_flush_l2())            # L2 flushed between timed steps (not timed)
        l0 = L.hgb_launch_count()
        step_ms.append(nb.step(1))                # CUDA events on the library's stream
        launches_timed += L.hgb_launch_count() - l0
    barrier()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = max_over_ranks(sum(step_ms) / len(step_ms))
    value = float(n) * n / (ms_per_step * 1e-3)
    ok, cnt = nbody_parity(nb, n, world, rank, seed=1)         # outside the timed region
    parity["nbody"] = {"ok": all_ok(ok), "targets": total(cnt), "of": "last timed step, every rank's copy"}

    # ---------------- nbody: end to end through the public call, host buffers ----------------
    e2e_steps = max(1, min(args.steps, 2))
    e2e_ms = []
    for i in range(e2e_steps + 1):
        barrier()
        t0 = time.perf_counter()
        if world == 1:
            # the drop-in for the reference's run(): H2D of the bodies, 1 step, D2H of the result
            H.nbody.run(pos_np, out_np, n, 1, 1)
        else:
            nb.upload(pos_np)
            nb.step(1)
            H._abi.check(L.hgb_nbody_download(nb._h, H._abi.dptr(out_np), None))
        barrier()
        if i:                                     # first pass is warm-up (allocations of the cached handle)
            e2e_ms.append((time.perf_counter() - t0) * 1e3)
    e2e_ms = max_over_ranks(sum(e2e_ms) / len(e2e_ms))
    nb.close()

    # like-for-like scaling base: when N=1 runs the pair-symmetric kernel, also time the standard
    # (one target per thread) kernel, same inputs, 1 warm-up + 1 step
    standard = None
    if sym:
        with H.nbody.NBody(n, ngpus=None, mode=H.NBODY_EXACT) as nb_std:
            nb_std.symmetric(0, 0).upload(pos_np)
            nb_std.step(1)
            std_kernel, std_par = nb_std.describe()
            barrier()
            H._abi.check(L.hgb_flush_l2())
            std_ms = max_over_ranks(nb_std.step(1))
        standard = {"kernel": std_kernel, "parallelism": std_par, "value": float(n) * n / (std_ms * 1e-3),
                    "unit": "interactions/s", "ms_per_step": std_ms, "steps": 1,
                    "fp64_instr_per_interaction": FP64_INSTR_PER_INTERACTION,
                    "note": "one target per thread, every interaction evaluated from both sides"}

    # ---------------- nbody: the other body sets of BASELINE.json / SURVEY 8(d) ----------------
    extra = {}
    if not args.skip_extra:
        for key, m, steps, gen, desc in (
                ("nbody_64k_plummer_100", 65536, 100, lambda: synth.plummer(65536, seed=1234),
                 "N=65536 Plummer sphere seed 1234, 100 device-resident steps (BASELINE configs[2])"),
                ("nbody_1m_plummer", N_BODIES, 2, lambda: synth.plummer(N_BODIES, seed=2),
                 "N=1048576 Plummer sphere seed 2 (SURVEY 8d item 4)")):
            if n != N_BODIES and m == N_BODIES:
                continue
            p0 = gen()
            with H.nbody.NBody(m, ngpus=None, mode=H.NBODY_EXACT) as nbx:
                nbx.upload(p0)
                nbx.step(1)
                nbx.upload(p0)
                barrier()
                H._abi.check(L.hgb_flush_l2())
                msx = max_over_ranks(nbx.step(steps))
                okx, cntx = nbody_parity(nbx, m, world, rank, seed=m)
                kernx = nbx.describe()[0]
                instrx = float(L.hgb_nbody_fp64_instr(nbx._h))
            parity[key] = {"ok": all_ok(okx), "targets": total(cntx), "of": f"step {steps}"}
            extra[key] = {"value": float(m) * m * steps / (msx * 1e-3), "unit": "interactions/s",
                          "ms_per_step": msx / steps, "steps": steps, "workload": desc,
                          "kernel": kernx, "fp64_instr_per_interaction": instrx}
            del p0

    # ---------------- convolution 16384^2 ------------------------------------------------------
    conv = {}
    if not args.skip_conv:
        cn = args.conv_size
        # EXACT mode (bit-identical, the parity path) first; then the opt-in FAST mode (one FMA per tap,
        # floating-point images only, 1e-12 relative) side by side.  FAST kernels are normalised so that ten
        # iterations keep the image in range.
        def unit(kern):
            return kern / np.abs(kern).sum()
        cases = [("3x3", 3, synth.kernel_dense(3), "float", False), ("7x7", 7, synth.kernel_dense(7), "float", False)]
        if not args.skip_extra:
            cases += [("5x5_kernel_txt", 5, synth.KERNEL_TXT.copy(), "float", False),
                      ("3x3_fast", 3, unit(synth.kernel_dense(3)), "float", True),
                      ("5x5_fast", 5, unit(synth.kernel_dense(5)), "float", True),
                      ("7x7_fast", 7, unit(synth.kernel_dense(7)), "float", True),
                      ("3x3_int_image", 3, synth.kernel_dense(3), "int", False)]
        pcie = pcie_probe(local_rank) if world == 1 else None
        dense, dense_kind = None, None
        for key, k, kern, kind, fast in cases:
            if kind != dense_kind:
                del dense
                dense = synth.image_float(cn, cn, seed=3) if kind == "float" else synth.image_int(cn, cn, seed=4)
                dense_kind = kind
            p = k // 2
            cv = H.convolution.Conv2D(cn, cn, p, ngpus=None)
            r0, rows = cv.band()
            band, band_t = pinned((rows, cn + 2 * p))           # pinned host buffers for the e2e leg
            got, got_t = pinned((rows, cn + 2 * p))
            if rows:
                band[...] = synth.add_padding(dense[r0:r0 + rows], p)
            if fast:
                cv.set_mode(H.CONV_FAST)
            cv.set_kernel(kern).upload(band)
            assert cv.mode_in_effect() == (H.CONV_FAST if fast else H.CONV_EXACT)
            cv.step(3)
            barrier()
            l0 = L.hgb_launch_count()
            ms = max_over_ranks(cv.step(CONV_TIMED))          # inputs (2 GiB) >> L2: no flush needed
            conv_launches = L.hgb_launch_count() - l0
            streamed = world == 1 and not fast
            if streamed:
                # the drop-in call (hgb_conv_run, state n-1 not requested): row chunks streamed over PCIe under
                # the iterations of other chunks.  In place, so the input is copied into `got` first (not timed);
                # one untimed call before it (chunk handle, staging buffers)
                cv.close()
                for timed in (False, True):
                    got[...] = band
                    barrier()
                    t0 = time.perf_counter()
                    H.convolution.run_flat(got, None, cn, cn, p, kern, CONV_ITERS, 1)
                    e2e = (time.perf_counter() - t0) * 1e3
            else:
                barrier()
                t0 = time.perf_counter()                      # e2e: H2D band, iterations, D2H band
                cv.upload(band)
                cv.step(CONV_ITERS)
                cv.download(out=got)
                barrier()
                e2e = max_over_ranks((time.perf_counter() - t0) * 1e3)
                cv.close()
            okc, cntc, errc = conv_parity(got, dense, r0, rows, cn, p, kern, CONV_ITERS, tol=1e-12 if fast else 0.0)
            parity[f"conv_{key}"] = {"ok": all_ok(okc), "rows": total(cntc),
                                     "of": f"{CONV_ITERS} iterations (the e2e leg's result), 3 bands of <= {PARITY_BAND} rows per rank incl. band seams",
                                     "criterion": "max-norm relative error <= 1e-12 (FAST mode)" if fast else "bit-identical",
                                     "worst_rel_err": max_over_ranks(errc)}
            del band, got, band_t, got_t
            gbps = cn * cn * BYTES_PER_PIXEL * CONV_TIMED / (ms * 1e-3) / 1e9
            peak = peaks.get("hbm_gbs", 6650.0)
            # 3x3 runs two iterations per pass (conv3_fused2_kernel): per LAUNCH the algorithmic bytes
            # are 2 x 16 B/pixel while DRAM moves the image once, so `frac` may exceed 1 and the
            # physical DRAM rate (`dram_frac`) is quoted beside it
            per_launch = 2 if k == 3 else 1
            traffic = None if fast else ncu_traffic("conv3_fused2_kernel" if k == 3 else f"conv_stream_kernel<{k},")
            launch_ms = ms / CONV_TIMED * per_launch
            image_desc = "uniform [0,1) doubles seed 3" if kind == "float" else "integer-valued doubles 0..255 seed 4"
            kern_desc = ("the shipped kernel.txt (5x5, 5 non-zero taps)" if key == "5x5_kernel_txt"
                         else f"dense {k}x{k} seed 5" + (" scaled to unit 1-norm" if fast else ""))
            entry = {"metric": "conv2d_achieved_GBps", "value": gbps, "unit": "GB/s",
                     "mode": "FAST (one FMA per tap; floating-point images only; not the parity path)" if fast else "EXACT (bit-identical)",
                     "ms_per_iteration": ms / CONV_TIMED, "iterations": CONV_TIMED, "gpu_launches": int(conv_launches),
                     "workload": f"{cn}x{cn} {image_desc}, {kern_desc}, row bands over {world} GPU(s)",
                     "e2e_GBps": cn * cn * BYTES_PER_PIXEL * CONV_ITERS / (e2e * 1e-3) / 1e9,
                     "e2e_call": ("hgb_conv_run(cur, prev=NULL): row chunks streamed, H2D / iterations / D2H on three streams"
                                  if streamed else "hgb_conv_upload + hgb_conv_step + hgb_conv_download of the rank's band, one after the other"),
                     "e2e_pcie_bound_GBps": (cn * cn * BYTES_PER_PIXEL * CONV_ITERS / (max(pcie["h2d_ms_per_gib"], pcie["d2h_ms_per_gib"])
                                             * (cn * (cn + 2 * p) * 8 / 2 ** 30) * 1e-3) / 1e9) if streamed else None,
                     "roofline": {"bound": "hbm", "achieved": gbps / world, "peak": peak,
                                  "unit": "GB/s", "frac": gbps / world / peak,
                                  "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "fallback 6650 (of fallback)",
                                  "kernel": ("conv3_fused2_kernel<2" if k == 3 else f"conv_stream_kernel<{k},256,2,3") + (",FAST>" if fast else ">"),
                                  "iterations_per_launch": per_launch,
                                  "algorithmic_bytes_per_launch": cn * cn * BYTES_PER_PIXEL * per_launch,
                                  "traffic": traffic,
                                  "traffic_note": "ncu dram bytes per launch at 16384^2 on 1 GPU (profiles/ncu_traffic.json)",
                                  "dram_frac": (traffic / (launch_ms * 1e-3) / 1e9 / peak) if (traffic and world == 1) else None,
                                  "per_gpu": True}}
            if k >= 5:   # 2*K^2 unfused FP64 ops per pixel make 5x5/7x7 FP64-pipe bound (FAST: K^2 FMAs)
                ops = (1.0 if fast else 2.0) * k * k * cn * cn * CONV_TIMED / (ms * 1e-3) / world
                issue = sms * 64.0 * (clocks["sm_mhz"] if clocks and clocks.get("sm_mhz") else peaks.get("sm_max_mhz", 1965.0)) * 1e6
                entry["roofline_fp64"] = {"bound": "fp64", "achieved": ops / 1e12, "peak": issue / 1e12,
                                          "unit": "T FP64-instr/s", "frac": ops / issue,
                                          "frac_of_probe": ops / 1e12 / (tf.value / 2),
                                          "peak_source": "clock-true issue rate: SMs x 64 lanes x SM clock sampled during the nbody steps",
                                          "note": "exact mode needs a DMUL and a DADD per tap (no FMA)"}
            conv[key] = entry
        if pcie:
            conv["pcie_probe"] = pcie
        del dense

    if rank == 0:
        sm_mhz = clocks["sm_mhz"] if clocks and clocks.get("sm_mhz") else None
        clock_true_tf = sms * 64 * 2 * sm_mhz * 1e6 / 1e12 if sm_mhz else None
        issue_rate = sms * 64 * sm_mhz * 1e6 if sm_mhz else tf.value / 2 * 1e12
        per_gpu = value / world
        all_parity_ok = all(v["ok"] for v in parity.values())
        line = {
            "metric": "nbody_fp64_interactions_per_s", "value": value, "unit": "interactions/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload(n),
                       "kernel": nb_kernel,
                       "parallelism": f"{world} GPU(s): {nb_parallelism}",
                       "l2": "512 MiB fill between timed steps (nbody); convolution inputs (2 GiB) exceed L2",
                       "timing": "CUDA events on the library stream per step, max over ranks",
                       # the other half of BASELINE's metric and the oracle verdict, where the driver's record keeps
                       # them verbatim (the full objects are the top-level "conv" and "parity_check" keys)
                       "conv_GBps": {k: round(v["value"], 1) for k, v in conv.items() if isinstance(v, dict) and "value" in v},
                       "conv_e2e_GBps": {k: round(v["e2e_GBps"], 1) for k, v in conv.items()
                                         if isinstance(v, dict) and v.get("e2e_GBps") is not None},
                       "parity_all_ok": all_parity_ok,
                       "parity_cases": {k: bool(v["ok"]) for k, v in parity.items()}},
            "wall_ms_per_step_incl_flush": wall_ms / args.steps,
            "e2e": {"value": float(n) * n / (e2e_ms * 1e-3), "unit": "interactions/s",
                    "h2d_bytes_per_step": 24 * n, "d2h_bytes_per_step": 24 * n,
                    "call": "hgb_nbody_run (drop-in for run()), pinned host arrays, 1 step per call" if world == 1
                            else "hgb_nbody_upload + hgb_nbody_step + hgb_nbody_download per rank"},
            "gpu_launches": int(launches_timed),
            "clocks": clocks,
            "parity_check": dict(parity, all_ok=all_parity_ok,
                                 checker="oracle/hg_oracle.c (C restatement pinned to the unmodified reference), bit-identical required"),
            "roofline": {"bound": "fp64", "achieved": per_gpu * FLOP_PER_INTERACTION / 1e12,
                         "peak": clock_true_tf if clock_true_tf else tf.value,
                         "unit": "TFLOP/s",
                         "frac": per_gpu * FLOP_PER_INTERACTION / 1e12 / (clock_true_tf if clock_true_tf else tf.value),
                         "kernel": nb_kernel,
                         "traffic": ncu_traffic("nbody_seg" if sym >= 4000 else "nbody_crow" if sym >= 3000 else "nbody_tri" if sym >= 2000 else "nbody_sym" if sym else "nbody_step_kernel<"),
                         "traffic_note": "ncu dram bytes per launch of this kernel on ONE GPU (profiles/ncu_traffic.json; nbody_crow_kernel captured at n = 2^18, the others at 2^20); positions (24 MiB) and hand-off words live in L2",
                         "per_gpu": True,
                         "peak_source": ("clock-true FP64 rate: SMs x 64 DFMA lanes x 2 flop x the SM clock sampled by nvidia-smi during "
                                         "the timed steps (MEASURED_PEAKS.json has no FP64 figure)") if clock_true_tf
                                        else "sustained DFMA probe (no clock samples)",
                         "peak_clock_true": clock_true_tf, "peak_probe": tf.value,
                         "probe": {"tflops": tf.value, "implied_sm_mhz": mhz.value, "clocks_during_probe": probe_clocks,
                                   "what": "~1 s of back-to-back dependent-free DFMA launches after 0.2 s of warm-up (hgb_measure_fp64_peak)"},
                         "frac_of_probe": per_gpu * FLOP_PER_INTERACTION / 1e12 / tf.value,
                         "fp64_instr_per_interaction": fp64_instr,
                         "fp64_pipe_util": per_gpu * fp64_instr / issue_rate,
                         "fp64_pipe_util_of_probe": per_gpu * fp64_instr / (tf.value / 2 * 1e12),
                         "fp64_pipe_util_note": (f"{fp64_instr:g} FP64-pipe instructions per interaction (SASS count, tools/sass_loop.py) x "
                                                 "interactions/s over the clock-true issue rate SMs x 64 x SM clock; comparable to ncu "
                                                 "sm__pipe_fp64_cycles_active (profiles/)"),
                         "sms": sms, "hbm_copy_probe_gbs": gbs.value},
            "conv": conv,
        }
        line.update(extra)
        if standard:
            line["nbody_standard_kernel"] = standard
        if world == 1 and not args.skip_cpu:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line), file=OUT, flush=True)
        if not all_parity_ok:
            print("PARITY MISMATCH: " + json.dumps(parity), file=sys.stderr, flush=True)
    barrier()
    L.hgb_finalize()
    if dist:
        dist.destroy_process_group()
    if not all(v["ok"] for v in parity.values()):
        sys.exit(3)


if __name__ == "__main__":
    main()
