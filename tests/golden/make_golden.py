"""Generate tests/golden/* from the UNMODIFIED reference (run in the dev container only).

Needs /root/reference and oracle/_ref (make -C oracle ref).  Outputs:
  conv_input_512.npz  the shipped 512x512 image `2d-convolution.input` as uint8
                      (`2d-convolution.txt` is the same image / 256, checked here)
  golden.json         results of the reference itself: raw doubles as hex, sha256
                      of whole result arrays, the CLI's printed lines
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"

from hg_c_benchmarks_b200 import synth  # noqa: E402
from oracle import ref  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def hexes(a):
    return [float(v).hex() for v in np.asarray(a).ravel()]


def main():
    ref.build(REF)
    g = {"nbody": {}, "conv": {}, "cli": {}}

    # ---- shipped image ---------------------------------------------------------
    img = ref.read_image(os.path.join(REF, "2d-convolution.input"))
    assert img.shape == (512, 512) and np.all(img == np.round(img)) and img.min() >= 0 and img.max() <= 255
    imgf = ref.read_image(os.path.join(REF, "2d-convolution.txt"))
    assert np.array_equal(imgf, img / 256.0)
    np.savez_compressed(os.path.join(HERE, "conv_input_512.npz"), image=img.astype(np.uint8))
    kernel_txt = ref.read_image(os.path.join(REF, "kernel.txt"))
    laplace_txt = ref.read_image(os.path.join(REF, "laplace.txt"))
    assert np.array_equal(kernel_txt, synth.KERNEL_TXT) and np.array_equal(laplace_txt, synth.LAPLACE_TXT)
    assert np.array_equal(ref.default_kernel(), synth.default_kernel())

    # ---- nbody: shipped initial condition, reference run() ---------------------------
    for n, iters in [(3, 0), (3, 1), (3, 2), (3, 3), (3, 10), (5, 10), (16, 7)]:
        a = ref.fill_planets(n)
        b = np.zeros_like(a)
        ref.nbody_run(a, b, iters, 1)
        g["nbody"][f"fill_planet_n{n}_it{iters}"] = {"planets": hexes(a), "buffer": hexes(b)}
    # seeded synthetic sets (hash of the whole arrays)
    for name, gen, n, iters in [("uniform", synth.uniform_cube, 257, 4), ("plummer", synth.plummer, 1000, 3),
                                ("uniform", synth.uniform_cube, 2049, 10), ("plummer", synth.plummer, 4096, 10)]:
        a = gen(n)
        b = np.zeros_like(a)
        ref.nbody_run(a, b, iters, 8)
        g["nbody"][f"{name}_n{n}_it{iters}"] = {"planets_sha256": sha(a), "buffer_sha256": sha(b),
                                                "planets_head": hexes(a[:2])}

    # ---- convolution: shipped inputs ---------------------------------------------------
    def conv_case(key, dense, kernel, iters):
        p = kernel.shape[1] // 2
        h, w = dense.shape
        cur = ref.add_padding(dense, p)
        prev = np.zeros_like(cur)
        ref.conv_run(cur, prev, w, h, p, kernel, iters, 8)
        cs = ref.conv_checksum(cur, w, h, p)
        g["conv"][key] = {"w": w, "h": h, "p": p, "iters": iters, "cur_sha256": sha(cur),
                          "prev_sha256": sha(prev), "checksum": float(cs).hex(),
                          "checksum_printed": "%f" % cs}

    for it in (1, 10, 100):
        conv_case(f"input_kernel_txt_n{it}", img, kernel_txt, it)
    for it in (1, 10, 100):
        conv_case(f"input_laplace_txt_n{it}", img, laplace_txt, it)
    conv_case("input_default5x5_n10", img, synth.default_kernel(), 10)
    conv_case("txt_kernel_txt_n10", imgf, kernel_txt, 10)
    # seeded synthetic images / dense kernels
    for w, h, k, it in [(33, 20, 7, 4), (300, 70, 3, 3), (513, 260, 5, 2), (1024, 768, 3, 5), (640, 480, 7, 3)]:
        conv_case(f"synth_float_{w}x{h}_k{k}_n{it}", synth.image_float(w, h), synth.kernel_dense(k), it)
    conv_case("synth_int_512x384_k3_n4", synth.image_int(512, 384), synth.kernel_dense(3), 4)
    conv_case("synth_int_512x384_k7_n2", synth.image_int(512, 384), synth.kernel_dense(7), 2)

    # ---- the reference's own gtest: run_laplace_neutralizes_image --------------------------
    im = np.array([[-1.5, 0, -1], [9, -2, 3], [2, 0, 5]], dtype=np.float64)
    cur = ref.add_padding(im, 1)
    prev = cur.copy()
    ref.conv_run(cur, prev, 3, 3, 1, synth.LAPLACE_TXT, 15, 1)
    g["conv"]["gtest_laplace_15"] = {"cur": hexes(cur), "prev": hexes(prev),
                                     "checksum": float(ref.conv_checksum(cur, 3, 3, 1)).hex()}

    # ---- the shipped CLI drivers (printed precision, as BASELINE.md quotes them) ------------
    for args in (["-s", 3, "-n", 10], ["-s", 3, "-n", 2], ["-s", 3, "-n", 3]):
        ref.cli("Nbody-OpenMP", args)
        g["cli"]["nbody " + " ".join(map(str, args))] = open(
            os.path.join(ROOT, "oracle/_ref/nbody.res")).read().splitlines()
    ref.cli("2D-Convolution", ["-f", os.path.join(REF, "2d-convolution.input"), "-k",
                               os.path.join(REF, "kernel.txt"), "-n", 10])
    g["cli"]["conv input kernel.txt n10"] = open(
        os.path.join(ROOT, "oracle/_ref/2d-convolution.res")).read().splitlines()

    json.dump(g, open(os.path.join(HERE, "golden.json"), "w"), indent=1, sort_keys=True)
    print("wrote golden.json with", sum(len(v) for v in g.values()), "entries")


if __name__ == "__main__":
    main()
