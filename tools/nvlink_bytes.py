"""NVLink bytes per nbody step of the fused compute+exchange kernels (VERDICT r2 missing #3): the driver's
per-link data counters (`nvidia-smi nvlink -gt d`) read before and after `steps` steps on every GPU of a
single-process multi-GPU handle, beside what the plan says must cross.

What the plan says: rows of RS = 32 targets are dealt over the GPUs in turns of `deal` rows; the last
row of a turn publishes its chain words -- the running column sums of accel() for every body below the row,
48 B per body (3 components x 16 B tagged word) -- into the next GPU's buffer.  Row I has 32 I bodies below
it, so one step moves  sum over turn ends I of 32 I x 48 B  ~= 1536 x T^2 / (2 deal)  bytes in total (T = n / 32
rows), 1/g of it out of each GPU, plus the finished sums (48 n bytes of tagged words) read by the g - 1 peers.

usage: python tools/nvlink_bytes.py [n] [g] [variant] [steps]   (variant 0: the default policy for that g)"""
import os
import re
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
g = int(sys.argv[2]) if len(sys.argv) > 2 else min(2, torch.cuda.device_count())
variant = int(sys.argv[3]) if len(sys.argv) > 3 else 0
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 4


def counters():
    """{gpu: (tx KiB, rx KiB)} summed over the links of each GPU"""
    out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d"], capture_output=True, text=True).stdout
    res, gpu = {}, None
    for line in out.splitlines():
        m = re.match(r"GPU (\d+):", line)
        if m:
            gpu = int(m.group(1))
            res[gpu] = [0, 0]
        m = re.search(r"Data Tx: (\d+) KiB", line)
        if m and gpu is not None:
            res[gpu][0] += int(m.group(1))
        m = re.search(r"Data Rx: (\d+) KiB", line)
        if m and gpu is not None:
            res[gpu][1] += int(m.group(1))
    if "N/A" in out and not any(a or b for a, b in res.values()):
        nv = nvml_counters()
        if nv:
            return nv, out
        return {}, out
    return res, out


def nvml_counters():
    """the same counters through NVML field values (KiB, all links of a GPU summed: scope id 0xffffffff), or {}"""
    try:
        import pynvml as nv
        nv.nvmlInit()
        res = {}
        for i in range(nv.nvmlDeviceGetCount()):
            h = nv.nvmlDeviceGetHandleByIndex(i)
            vals = nv.nvmlDeviceGetFieldValues(h, [(nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, 0xffffffff),
                                                   (nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, 0xffffffff)])
            if any(v.nvmlReturn != 0 for v in vals):
                return {}
            res[i] = [int(vals[0].value.ullVal), int(vals[1].value.ullVal)]
        return res
    except Exception as e:   # noqa: BLE001 -- a diagnostic tool: say why and go on with the plan numbers
        print(f"NVML field values unavailable: {e!r}")
        return {}


pos = synth.uniform_cube(n, seed=1)
with H.nbody.NBody(n, ngpus=g, exchange=H._abi.EXCHANGE_P2P) as nb:
    if variant:
        nb.symmetric(variant, 0)
    nb.upload(pos)
    nb.step(1)
    what = nb.describe()
    before, raw = counters()
    ms = nb.step(steps) / steps
    after, _ = counters()
print(f"n={n} g={g} steps={steps} {ms:.2f} ms/step  [{what}]")
if not before:
    print("the driver exposes no NVLink data counters on this box (nvidia-smi nvlink -gt d below; NVML field values likewise):\n" + raw[:300])
    after = {}
# rows per turn as tri_default_deal() has them (no hgb_nbody_set_deal here); rows of 64 targets for the segmented kernel
kernel = what[0]
deal = 16 if "nbody_tri" in kernel else 8 if ("nbody_crow" in kernel or "nbody_seg" in kernel) else None
RS = 64 if "nbody_seg" in kernel else 32
T = (n + RS - 1) // RS
for gpu in sorted(after):
    tx = (after[gpu][0] - before[gpu][0]) * 1024 / steps
    rx = (after[gpu][1] - before[gpu][1]) * 1024 / steps
    print(f"GPU {gpu}: tx {tx / 1e9:8.3f} GB/step ({tx / (ms * 1e-3) / 1e9:7.1f} GB/s)   rx {rx / 1e9:8.3f} GB/step ({rx / (ms * 1e-3) / 1e9:7.1f} GB/s)")
if deal:
    total = 48.0 * RS * T * T / (2 * deal)
    print(f"plan: chain words {total / 1e9:.3f} GB/step in total = {total / g / 1e9:.3f} GB out of each GPU (turns of {deal} rows of {RS}), "
          f"finished sums {48.0 * n * (g - 1) / 1e9:.3f} GB read from the last row's GPU")
