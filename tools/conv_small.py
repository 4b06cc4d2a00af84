"""Dev probe: single-GPU convolution time for band-sized images (what one rank of 8 computes)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hg_c_benchmarks_b200 as H
from hg_c_benchmarks_b200 import synth
W = 16384
for k in [int(x) for x in os.environ.get("KS", "3,7").split(",")]:
    p = k // 2
    for rows in [int(x) for x in os.environ.get("ROWS", "512,2048,4096,8192").split(",")]:
        img = synth.add_padding(synth.image_float(W, rows), p)
        for chunk in [int(x) for x in os.environ.get("CHUNKS", "32,64,128,256").split(",")]:
            cv = H.convolution.Conv2D(W, rows, p)
            cv.set_kernel(synth.kernel_dense(k)).tune(0, chunk).upload(img)
            cv.step(5)
            ms = cv.step(50) / 50
            print(f"k={k} rows={rows} chunk={chunk}: {ms*1e3:.1f} us/iter  {W*rows*16/(ms*1e-3)/1e9:.0f} GB/s", flush=True)
            cv.close()
