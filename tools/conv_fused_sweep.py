"""3x3 at 16384^2 on one GPU: plain pass vs two-iterations-per-pass kernel over chunk heights."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hg_c_benchmarks_b200 as H  # noqa: E402
from hg_c_benchmarks_b200 import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
img = synth.image_float(n, n, seed=3)
with H.convolution.Conv2D(n, n, 1) as cv:
    cv.set_kernel(synth.kernel_dense(3)).upload_dense(img)
    ref = None
    for variant, band in [(4, 0), (8, 64), (0, 32), (0, 64), (0, 128), (16, 16), (16, 32), (16, 48), (16, 64), (16, 96), (16, 128), (16, 256)]:
        cv.tune(variant, band).upload_dense(img)
        cv.step(5)
        chk = cv.checksum() if n <= 4096 else None
        ms = cv.step(101) / 101
        print(f"variant={variant} band={band}: {ms * 1e3:8.1f} us/iter  {n * n * 16 / ms / 1e6:8.1f} GB/s  checksum={chk}", flush=True)
H.lib().hgb_finalize()
