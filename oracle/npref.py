"""numpy restatement of the hot path (small sizes).  TEST INFRASTRUCTURE ONLY.

numpy float64 elementwise + - * / sqrt are single IEEE-754 operations (never
fused), so keeping the reference's association and summation ORDER makes this
bit-identical to the C code.  Vectorisation is across targets / pixels only;
the j loop (nbody) and the tap loop (convolution) stay sequential.
"""
import numpy as np

G = 9.8      # src/nbody/nbody-run.h:12
EPS = 0.005  # src/nbody/nbody-run.h:13


def nbody_step(pos):
    """accel() for every target (src/nbody/nbody-run.c:6-25)."""
    n = pos.shape[0]
    px, py, pz = pos[:, 0].copy(), pos[:, 1].copy(), pos[:, 2].copy()
    ax = np.zeros(n)
    ay = np.zeros(n)
    az = np.zeros(n)
    for j in range(n):
        dx = px[j] - px
        dy = py[j] - py
        dz = pz[j] - pz
        s = ((dx * dx + dy * dy) + dz * dz) + EPS
        f = 1.0 / np.sqrt((s * s) * s)
        ax = ax + dx * f
        ay = ay + dy * f
        az = az + dz * f
    return np.stack([G * ax, G * ay, G * az], axis=1)


def nbody_run(planets, buffer, iterations):
    """run() incl. the local-pointer-swap parity (src/nbody/nbody-run.c:27-37)."""
    src, dst = planets, buffer
    for _ in range(iterations):
        dst[...] = nbody_step(src)
        src, dst = dst, src


def conv_step(inp, w, h, p, kernel):
    """apply_kernel_to_padded_image + update_borders on (h, w+2p) rows
    (src/convolution/convolution-run.c:52-81, convolution-util.c:281-294)."""
    kh, kw = kernel.shape
    out = np.empty_like(inp)
    val = np.zeros((h, w))
    rows = np.arange(h)
    for ky in range(kh):
        src_rows = np.clip(rows + ky - p, 0, h - 1)   # aliased pad rows == clamp
        for kx in range(kw):
            val = val + kernel[ky, kx] * inp[src_rows, kx:kx + w]
    out[:, p:p + w] = val
    for x in range(p):
        out[:, x] = out[:, p]
        out[:, w + 2 * p - 1 - x] = out[:, p + w - 1]
    return out


def conv_run(cur, prev, w, h, p, kernel, iterations):
    a, b = cur, prev
    for _ in range(iterations):
        b[...] = conv_step(a, w, h, p, kernel)
        a, b = b, a
    if iterations & 1:
        t = cur.copy()
        cur[...] = prev
        prev[...] = t


def checksum(flat, w, h, p):
    """get_checksum: sequential row-major sum (convolution-util.c:63-71)."""
    val = 0.0
    for v in flat[:, p:p + w].ravel().tolist():
        val += v
    return val
