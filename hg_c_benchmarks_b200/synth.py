"""Deterministic synthetic inputs shared by the CPU checker and the GPU path.

SURVEY.md §8(d): seeded uniform cube / Plummer sphere body sets and random images.
The generator is splitmix64 (counter based, so arrays of any size are reproducible
and independent of chunking); doubles are the top 53 bits / 2^53, i.e. exact
dyadic rationals in [0, 1).
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(seed, count, offset=0):
    """count uint64 values of the splitmix64 stream `seed`, starting at index `offset`."""
    with np.errstate(over="ignore"):
        idx = np.arange(offset + 1, offset + count + 1, dtype=np.uint64)
        z = np.uint64(seed) + idx * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def uniform01(seed, count, offset=0):
    return (splitmix64(seed, count, offset) >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def uniform_cube(n, seed=1):
    """n bodies uniform in [0,1)^3, AoS (n,3) float64 == Float3D[n]."""
    return uniform01(seed, 3 * n).reshape(n, 3).copy()


def plummer(n, seed=2, a=1.0):
    """Plummer sphere, scale radius a; mass fraction m in (0.0005, 0.9995), r = a/sqrt(m^(-2/3)-1)."""
    u = uniform01(seed, 3 * n).reshape(n, 3)
    m = 0.0005 + 0.999 * u[:, 0]
    r = a / np.sqrt(m ** (-2.0 / 3.0) - 1.0)
    cos_t = 2.0 * u[:, 1] - 1.0
    sin_t = np.sqrt(np.maximum(0.0, 1.0 - cos_t * cos_t))
    phi = 2.0 * np.pi * u[:, 2]
    pos = np.empty((n, 3), dtype=np.float64)
    pos[:, 0] = r * sin_t * np.cos(phi)
    pos[:, 1] = r * sin_t * np.sin(phi)
    pos[:, 2] = r * cos_t
    return pos


def fill_planets(n):
    """The shipped initial condition (i, 0.2 i, 30 i)  (reference src/nbody/nbody-util.c:92-96)."""
    i = np.arange(n, dtype=np.float64)
    return np.stack([i * 1.0, i * 0.2, i * 30.0], axis=1)


def image_float(w, h, seed=3):
    """dense (h, w) image of uniform [0,1) doubles."""
    return uniform01(seed, w * h).reshape(h, w).copy()


def image_int(w, h, seed=4):
    """dense (h, w) image of integer-valued doubles 0..255."""
    return (splitmix64(seed, w * h) >> np.uint64(56)).astype(np.float64).reshape(h, w).copy()


def kernel_dense(k, seed=5):
    """dense k x k taps uniform in [-1, 1)."""
    return (2.0 * uniform01(seed, k * k) - 1.0).reshape(k, k).copy()


KERNEL_TXT = np.array([[0, 0, 0, 0, 0],
                       [0, 0, 0.125, 0, 0],
                       [0, 0.125, 0.5, 0.125, 0],
                       [0, 0, 0.125, 0, 0],
                       [0, 0, 0, 0, 0]], dtype=np.float64)       # shipped kernel.txt
LAPLACE_TXT = np.array([[0, 0.25, 0], [0.25, -1, 0.25], [0, 0.25, 0]], dtype=np.float64)  # laplace.txt


def default_kernel():
    """get_default_kernel (reference src/convolution/convolution-util.c:135-151)."""
    k = np.where((np.add.outer(np.arange(5), np.arange(5)) % 2) == 0, 2.0, -2.0)
    k[2, 2] = -1.0
    return k


def add_padding(dense, p):
    """Dense (h,w) -> pitched rows (h, w+2p) with clamp-to-edge pad columns
    (reference add_padding, src/convolution/convolution-util.c:258-269)."""
    dense = np.ascontiguousarray(dense, dtype=np.float64)
    if p == 0:
        return dense.copy()
    return np.ascontiguousarray(np.pad(dense, ((0, 0), (p, p)), mode="edge"))


def velocities(n, seed=6, scale=0.5):
    """(n,3) velocities uniform in [-scale, scale) for the integrator extension."""
    return ((2.0 * uniform01(seed, 3 * n) - 1.0) * scale).reshape(n, 3).copy()


def masses(n, seed=7):
    """(n,) masses uniform in [0.5, 1.5) / n  (total mass ~ 1)."""
    return (0.5 + uniform01(seed, n)) / float(n)
