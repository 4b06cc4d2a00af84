"""Instruction mix of the innermost hot loop of a kernel, from `cuobjdump -sass` (no GPU needed).
usage: python tools/sass_loop.py <object-or-so> <mangled-kernel-name> [min_fp64_in_loop]"""
import re
import subprocess
import sys
from collections import Counter


def main(obj, fun, min_fp=8):
    sass = subprocess.run(["cuobjdump", "-sass", "-fun", fun, obj], capture_output=True, text=True).stdout
    ins = []
    for l in sass.splitlines():
        m = re.search(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2)))
    addr = {a: i for i, (a, _) in enumerate(ins)}
    fp = lambda x: re.match(r"(@\S+\s+)?D(FMA|MUL|ADD)", x) is not None
    loops = []
    for i, (a, t) in enumerate(ins):
        m = re.search(r"BRA\S*\s+.*0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a and int(m.group(1), 16) in addr:
            body = ins[addr[int(m.group(1), 16)]:i + 1]
            n = sum(1 for _, x in body if fp(x))
            if n >= min_fp:
                loops.append((len(body), n, int(m.group(1), 16), a, body))
    loops.sort()
    for size, n, lo, hi, body in loops[:3]:
        c = Counter(re.sub(r"^@\S+\s+", "", x).split()[0].split(".")[0] for _, x in body)
        print(f"loop {lo:#x}..{hi:#x}: {size} instructions, {n} FP64-pipe: {c.most_common()}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 8)
