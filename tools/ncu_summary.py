"""Summarise an .ncu-rep (read here, no GPU needed) into a small text file for profiles/.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/out.txt"""
import csv
import os
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "sm__cycles_elapsed.avg.per_second", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "dram__bytes_write.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__warps_eligible.avg.per_cycle_active",
]
STALL = "smsp__average_warps_issue_stalled_"


def traffic_json(reps, out):
    """per-kernel DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, mean over the captured
    launches) -> profiles/ncu_traffic.json, which bench.py quotes as roofline.traffic"""
    import json
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    acc = {}
    for rep in reps:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(l for l in raw.splitlines() if l.startswith('"')))
        hdr, units = rows[0], rows[1]
        ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
        for r in rows[2:]:
            name = r[ik].split("(")[0].replace("void ", "").replace("hgb::", "").strip()
            acc.setdefault(name, []).append(float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]])
    res = {k: {"dram_bytes_per_launch": sum(v) / len(v), "launches": len(v)} for k, v in acc.items()}
    res["_source"] = "ncu --set full --clock-control none: " + ", ".join(os.path.basename(r) for r in reps)
    with open(out, "w") as f:
        json.dump(res, f, indent=1, sort_keys=True)
        f.write("\n")


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(l for l in raw.splitlines() if l.startswith('"')))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        f.write(f"# ncu --set full --clock-control none, summarised from {rep.split('/')[-1]} by tools/ncu_summary.py\n")
        for n, r in enumerate(rows[2:]):
            name = r[hdr.index("Kernel Name")]
            f.write(f"\n== launch {n}: {name}\n")
            for i, h in enumerate(hdr):
                if h in KEEP:
                    f.write(f"{h:70s} {r[i]} {units[i]}\n")
            stalls = []
            for i, h in enumerate(hdr):
                if h.startswith(STALL) and h.endswith("_per_issue_active.ratio") and "not_issued" not in h:
                    try:
                        stalls.append((float(r[i]), h[len(STALL):-len("_per_issue_active.ratio")]))
                    except ValueError:
                        pass
            stalls.sort(reverse=True)
            f.write("stall reasons (warps per issue-active cycle): " +
                    ", ".join(f"{k}={v:.2f}" for v, k in stalls[:7]) + "\n")
            try:
                rd = float(r[hdr.index("dram__bytes_read.sum")])
                wr = float(r[hdr.index("dram__bytes_write.sum")])
                ur, uw = units[hdr.index("dram__bytes_read.sum")], units[hdr.index("dram__bytes_write.sum")]
                scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
                f.write(f"dram traffic per launch (read+write)                                   "
                        f"{(rd * scale[ur] + wr * scale[uw]) / 1e9:.4f} GB\n")
            except Exception:
                pass


if __name__ == "__main__":
    if sys.argv[1] == "--traffic":      # python tools/ncu_summary.py --traffic out.json rep1 rep2 ...
        traffic_json(sys.argv[3:], sys.argv[2])
    else:
        main(sys.argv[1], sys.argv[2])
