"""Kernel shares of a step from an ncu launch list (--metrics gpu__time_duration.sum --csv).
usage: python tools/launch_shares.py gpurun_out/launches.csv profiles/out.txt "<command that was profiled>" """
import csv
import sys
from collections import defaultdict


def main(src, out, cmd):
    rows = list(csv.reader(l for l in open(src) if l.startswith('"')))
    hdr = rows[0]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows[1:]:
        name = r[ik].split("(")[0].strip()
        tot[name] += float(r[iv].replace(",", "")) * scale[r[iu]]
        cnt[name] += 1
    total = sum(tot.values())
    with open(out, "w") as f:
        f.write(f"# kernel share of `{cmd}` under ncu --metrics gpu__time_duration.sum --clock-control none "
                "(serialised, cold cache: compare SHARES)\n")
        f.write(f"# total {total:.1f} ms over {sum(cnt.values())} launches\n")
        for name in sorted(tot, key=tot.get, reverse=True):
            f.write(f"{tot[name]:12.3f} ms {100 * tot[name] / total:6.2f}% x{cnt[name]:3d}  {name[:90]}\n")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "python bench.py --steps 1 --warmup 1")
