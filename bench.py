#!/usr/bin/env python
"""bench.py -- headline benchmark of the hg-c-benchmarks hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One "step" = one all-pairs FP64 nbody step over N = 2^20 bodies (BASELINE.json configs[3]),
EXACT mode (bit-identical to the reference).  Rank 0 prints ONE JSON line.  The second
benchmark kernel (16384^2 2D convolution, 3x3 and 7x7, configs[4]) is reported in the
same line under "conv".  `--impl reference` times the unmodified reference (oracle/_ref,
else the oracle port) on the host cores for the same metric.
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
CONV_ITERS = 10          # the config's iteration count (used by the end-to-end leg)
CONV_TIMED = 100         # device-resident timing runs 10 x 10 iterations back to back: at 8 GPUs ten
                         # iterations last ~1 ms, comparable to the host skew between ranks after a barrier
FLOP_PER_INTERACTION = 19          # 3 sub, 3 mul+3 add (+EPS), 2 mul, sqrt, div, 3 mul+3 add
FP64_INSTR_PER_INTERACTION = 29    # exact mode: 17 plain + 8 (IEEE sqrt) + 4 (IEEE reciprocal) FP64-pipe ops (tools/sass_loop.py)
FP64_INSTR_PER_INTERACTION_SYM = 16  # pair-symmetric kernel: 26 + 3 + 3 FP64-pipe ops per PAIR = 2 interactions
BYTES_PER_PIXEL = 16               # 8 B read + 8 B written per inner pixel per iteration


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


def pinned(shape):
    """numpy array backed by pinned host memory (the e2e legs copy from/to pinned buffers)"""
    import torch
    t = torch.empty(shape, dtype=torch.float64, pin_memory=True)
    return t.numpy(), t


# ---------------------------------------------------------------------------------------------
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

    # calibrate so one step is a few seconds of wall time; where the -O3 + profile-feedback build
    # of the reference (what pgo.sh intended, oracle/Makefile ref-pgo) exists, the faster of the
    # two builds on this host is the one timed
    cal = synth.uniform_cube(8192, seed=1)
    buf = np.zeros_like(cal)
    builds = {}
    for variant in (("", "_pgo") if use_ref else ("",)):
        if use_ref:
            if not ref.available(variant):
                continue
            ref.use(variant)
        step(cal, buf)
        t0 = time.perf_counter()
        step(cal, buf)
        builds[variant] = 8192.0 ** 2 / (time.perf_counter() - t0)
    best = max(builds, key=builds.get)
    if use_ref:
        ref.use(best)
    rate = builds[best]
    build_name = {"": "shipped flags (-std=c11 -O2 -fopenmp)", "_pgo": "-O3 + gcc profile feedback (pgo.sh's intent)"}[best]
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
    line = {"impl": "reference", "metric": "nbody_fp64_interactions_per_s", "value": value,
            "unit": "interactions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload(N_BODIES),
                       "sample": "CPU arm: a bounded sample of the workload per step, see cpu_baseline.sample"},
            "cpu_baseline": {"value": value, "unit": "interactions/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "interactions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=OUT, flush=True)


# ---------------------------------------------------------------------------------------------
def cpu_baseline():
    """bounded CPU sample on rank 0 at N=1: the unmodified reference on all host cores"""
    from hg_c_benchmarks_b200 import synth
    from oracle import port, ref
    cores = port.ncpu()
    use_ref = ref.available()
    run = (lambda a, b: ref.nbody_run(a, b, 1, cores)) if use_ref else (lambda a, b: port.nbody_run(a, b, 1, cores))
    cal = synth.uniform_cube(8192, seed=1)
    buf = np.zeros_like(cal)
    run(cal, buf)
    t0 = time.perf_counter()
    run(cal, buf)
    rate = 8192.0 ** 2 / (time.perf_counter() - t0)
    n = int(min(262144, max(16384, (rate * 12.0) ** 0.5))) // 1024 * 1024
    pos = synth.uniform_cube(N_BODIES, seed=1)[:n].copy()
    buf = np.zeros_like(pos)
    t0 = time.perf_counter()
    run(pos, buf)
    dt = time.perf_counter() - t0
    out = {"value": float(n) * n / dt, "unit": "interactions/s", "cores": cores,
           "kind": "reference" if use_ref else "port",
           "sample": f"first {n} of the {N_BODIES} bodies, one step ({n}^2 interactions, {dt:.1f} s), "
                     f"reference run() with number_of_processes={cores}",
           "build": "shipped effective flags: -std=c11 -fcommon -O2 -fopenmp (CMakeLists.txt:40,45)"}
    if use_ref and ref.available("_pgo"):
        # the reference's pgo.sh is inert as shipped; its intent (-O3 + profile feedback) is built by
        # oracle/Makefile ref-pgo from the same unmodified sources and timed on a smaller sample
        ref.use("_pgo")
        m = max(8192, n // 2 // 1024 * 1024)
        q = pos[:m].copy()
        qb = np.zeros_like(q)
        t0 = time.perf_counter()
        ref.nbody_run(q, qb, 1, cores)
        dq = time.perf_counter() - t0
        ref.use("")
        out["pgo"] = {"value": float(m) * m / dq, "unit": "interactions/s",
                      "build": "-O3 + gcc -fprofile-use trained by the reference CLI (-s 4096 -n 2)",
                      "sample": f"first {m} bodies, one step ({dq:.1f} s)"}
    # convolution: 4096^2, 3x3 dense, 2 iterations
    w = 4096
    img = synth.add_padding(synth.image_float(w, w, seed=3), 1)
    prev = np.zeros_like(img)
    k = synth.kernel_dense(3)
    conv = (lambda: ref.conv_run(img, prev, w, w, 1, k, 2, cores)) if use_ref else (lambda: port.conv_run(img, prev, w, w, 1, k, 2, cores))
    t0 = time.perf_counter()
    conv()
    dtc = time.perf_counter() - t0
    out["conv"] = {"value": w * w * 2 * BYTES_PER_PIXEL / dtc / 1e9, "unit": "GB/s",
                   "sample": f"{w}x{w} 3x3 dense, 2 iterations, run_on_padded_image with {cores} threads"}
    return out


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

    peaks = measured_peaks()
    tf, mhz, gbs = C.c_double(), C.c_double(), C.c_double()
    H._abi.check(L.hgb_measure_fp64_peak(C.byref(tf), C.byref(mhz)))
    H._abi.check(L.hgb_measure_hbm_copy(C.byref(gbs)))

    n = args.bodies
    pos_np, pos_t = pinned((n, 3))
    pos_np[...] = synth.uniform_cube(n, seed=1)
    out_np, out_t = pinned((n, 3))

    # ---------------- nbody: device-resident steps -------------------------------------------
    nb = H.nbody.NBody(n, ngpus=None, mode=H.NBODY_EXACT)
    nb.upload(pos_np)
    sym = int(L.hgb_nbody_symmetric(nb._h))      # pair-symmetric variant in use (0: standard kernel)
    fp64_instr = FP64_INSTR_PER_INTERACTION_SYM if sym else FP64_INSTR_PER_INTERACTION
    nb_kernel = f"nbody_sym_kernel<{sym // 100},{sym // 10 % 10},{sym % 10}>" if sym else "nbody_step_kernel<128,2,24,false>"
    for _ in range(args.warmup):
        nb.step(1)
    sampler = ClockSampler(local_rank)
    launches0 = L.hgb_launch_count()
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
    # (one target per thread) kernel that the multi-GPU runs use, same inputs, 1 warm-up + 1 step
    standard = None
    if sym and world == 1:
        with H.nbody.NBody(n, ngpus=None, mode=H.NBODY_EXACT) as nb_std:
            nb_std.symmetric(0, 0).upload(pos_np)
            nb_std.step(1)
            H._abi.check(L.hgb_flush_l2())
            std_ms = nb_std.step(1)
        standard = {"kernel": "nbody_step_kernel<128,2,24,false>", "value": float(n) * n / (std_ms * 1e-3),
                    "unit": "interactions/s", "ms_per_step": std_ms, "steps": 1,
                    "note": "the kernel every rank runs at N > 1 (targets sharded); scaling of that kernel is value(N) / this"}

    # ---------------- convolution 16384^2, 3x3 and 7x7 ----------------------------------------
    conv = {}
    if not args.skip_conv:
        cn = args.conv_size
        dense = synth.image_float(cn, cn, seed=3)
        for k in (3, 7):
            p = k // 2
            cv = H.convolution.Conv2D(cn, cn, p, ngpus=None)
            r0, rows = cv.band()
            band, band_t = pinned((rows, cn + 2 * p))           # pinned host buffers for the e2e leg
            got, got_t = pinned((rows, cn + 2 * p))
            if rows:
                band[...] = synth.add_padding(dense[r0:r0 + rows], p)
            cv.set_kernel(synth.kernel_dense(k)).upload(band)
            cv.step(3)
            barrier()
            ms = max_over_ranks(cv.step(CONV_TIMED))          # inputs (2 GiB) >> L2: no flush needed
            barrier()
            t0 = time.perf_counter()                          # e2e: H2D band, iterations, D2H band
            cv.upload(band)
            cv.step(CONV_ITERS)
            cv.download(out=got)
            barrier()
            e2e = max_over_ranks((time.perf_counter() - t0) * 1e3)
            cv.close()
            del band, got, band_t, got_t
            gbps = cn * cn * BYTES_PER_PIXEL * CONV_TIMED / (ms * 1e-3) / 1e9
            peak = peaks.get("hbm_gbs", 6650.0)
            # 3x3 runs two iterations per pass (conv3_fused2_kernel): per LAUNCH the algorithmic bytes
            # are 2 x 16 B/pixel while DRAM moves the image once, so `frac` may exceed 1 and the
            # physical DRAM rate (`dram_frac`) is quoted beside it
            per_launch = 2 if k == 3 else 1
            traffic = ncu_traffic("conv3_fused2_kernel" if k == 3 else f"conv_stream_kernel<{k},")
            launch_ms = ms / CONV_TIMED * per_launch
            entry = {"metric": "conv2d_achieved_GBps", "value": gbps, "unit": "GB/s",
                     "ms_per_iteration": ms / CONV_TIMED, "iterations": CONV_TIMED,
                     "workload": f"{cn}x{cn} uniform [0,1) doubles seed 3, dense {k}x{k} seed 5, row bands over {world} GPU(s)",
                     "e2e_GBps": cn * cn * BYTES_PER_PIXEL * CONV_ITERS / (e2e * 1e-3) / 1e9,
                     "roofline": {"bound": "hbm", "achieved": gbps / world, "peak": peak,
                                  "unit": "GB/s", "frac": gbps / world / peak,
                                  "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "fallback 6650 (of fallback)",
                                  "kernel": "conv3_fused2_kernel<2>" if k == 3 else f"conv_stream_kernel<{k},256,2,3>",
                                  "iterations_per_launch": per_launch,
                                  "algorithmic_bytes_per_launch": cn * cn * BYTES_PER_PIXEL * per_launch,
                                  "traffic": traffic,
                                  "traffic_note": "ncu dram bytes per launch at 16384^2 on 1 GPU (profiles/ncu_traffic.json)",
                                  "dram_frac": (traffic / (launch_ms * 1e-3) / 1e9 / peak) if (traffic and world == 1) else None,
                                  "per_gpu": True}}
            if k >= 5:   # 2*K^2 unfused FP64 ops per pixel make 5x5/7x7 FP64-pipe bound
                ops = 2.0 * k * k * cn * cn * CONV_TIMED / (ms * 1e-3) / world
                entry["roofline_fp64"] = {"bound": "fp64", "achieved": ops / 1e12, "peak": tf.value / 2,
                                          "unit": "T FP64-instr/s", "frac": ops / 1e12 / (tf.value / 2),
                                          "note": "exact mode needs a DMUL and a DADD per tap (no FMA)"}
            conv[f"{k}x{k}"] = entry
        del dense

    if rank == 0:
        line = {
            "metric": "nbody_fp64_interactions_per_s", "value": value, "unit": "interactions/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload(n),
                       "kernel": nb_kernel,
                       "parallelism": f"targets sharded over {world} GPU(s), ncclAllGather of positions per step",
                       "l2": "512 MiB fill between timed steps (nbody); convolution inputs (2 GiB) exceed L2",
                       "timing": "CUDA events on the library stream per step, max over ranks"},
            "wall_ms_per_step_incl_flush": wall_ms / args.steps,
            "e2e": {"value": float(n) * n / (e2e_ms * 1e-3), "unit": "interactions/s",
                    "h2d_bytes_per_step": 24 * n, "d2h_bytes_per_step": 24 * n,
                    "call": "hgb_nbody_run (drop-in for run()), pinned host arrays, 1 step per call" if world == 1
                            else "hgb_nbody_upload + hgb_nbody_step + hgb_nbody_download per rank"},
            "gpu_launches": int(launches_timed),
            "clocks": clocks,
            "roofline": {"bound": "fp64", "achieved": value / world * FLOP_PER_INTERACTION / 1e12, "peak": tf.value,
                         "unit": "TFLOP/s", "frac": value / world * FLOP_PER_INTERACTION / 1e12 / tf.value,
                         "kernel": nb_kernel,
                         "traffic": ncu_traffic("nbody_sym_kernel<" if sym else "nbody_step_kernel<"),
                         "traffic_note": ("ncu dram bytes per launch (profiles/ncu_traffic.json): positions (24 MiB) stay in L2; the bytes are "
                                          "evictions of the tagged chain words handed between workers, ~1 GB/s, <0.02 % of HBM bandwidth")
                                         if sym else "ncu dram bytes per launch (profiles/ncu_traffic.json): the 24 MiB of positions stay in L2",
                         "per_gpu": True,
                         "peak_source": "in-run dependent-free DFMA probe (MEASURED_PEAKS.json has no FP64 figure)",
                         "fp64_pipe_util": value / world * fp64_instr / (tf.value / 2 * 1e12),
                         "fp64_pipe_util_note": f"{fp64_instr} FP64-pipe instructions per interaction in exact mode (SASS count) vs measured DFMA issue rate",
                         "probe_sm_mhz": mhz.value, "hbm_copy_probe_gbs": gbs.value},
            "conv": conv,
        }
        if standard:
            line["nbody_standard_kernel"] = standard
        if world == 1 and not args.skip_cpu:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line), file=OUT, flush=True)
    barrier()
    L.hgb_finalize()
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
