"""The segmented first-generation schedule (csrc/nbody_seg.cuh) checked on the CPU: tests/seg_model.py mirrors the
kernel's item logic (stream limits per item, skew fill / drain, chunk open / close per item, row sums between items,
ticket plan with 64-target blocks, blocks dealt over GPUs in turns); its result must be bit-identical to the oracle's
in-order sum (accel(), reference src/nbody/nbody-run.c:17-25)."""
import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth
from seg_model import SegModel


@pytest.mark.parametrize("n,seg,rows,weight,world,deal", [(1, 32, 1, 1, 1, None), (31, 32, 2, 1, 1, None), (64, 32, 1, 1, 1, None),
                                                         (65, 64, 2, 1, 1, None), (129, 32, 1, 2, 1, None), (200, 64, 3, 2, 1, None),
                                                         (333, 96, 2, 5, 1, None), (300, 8192, 16, 16, 1, None), (450, 64, 2, 1, 2, None),
                                                         (450, 32, 1, 4, 3, None), (520, 160, 3, 2, 4, None), (600, 96, 1, 7, 2, 3),
                                                         (700, 128, 2, 16, 4, 6), (641, 160, 1, 1, 8, 1)])
def test_schedule_reproduces_the_in_order_sum(n, seg, rows, weight, world, deal):
    from oracle import port
    pos = synth.uniform_cube(n, seed=n)
    if n >= 64:
        pos[5] = pos[40]                    # duplicate body: d = +0 pairs
        pos[63, 0] = pos[3, 0]
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    m = SegModel([tuple(map(float, p)) for p in pos], seg, rows, weight, world, deal)
    got = np.array(m.step(), dtype=np.float64).reshape(n, 3)
    assert np.array_equal(got, want)
    assert m.items == sum(64 * w // seg + 1 for w in range(m.T))
