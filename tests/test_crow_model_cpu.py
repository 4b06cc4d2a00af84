"""The third-generation pair-symmetric nbody schedule (csrc/nbody_crow.cuh: a CTA of W warps per row, the
row's ordered sums rippling through the warps as tokens) checked on the CPU: tests/crow_model.py mirrors the
kernel's inner-loop index arithmetic (minitile -> warp, per-warp source rings, chain-word pointers, word
registers, peeled / paired / trailing minitiles, tokens); its result must be bit-identical to the oracle's
in-order sum (accel(), reference src/nbody/nbody-run.c:17-25) for ragged sizes, segment sizes whose minitile
count per warp is 1, 2, 3, even, odd and not a multiple of 4, 4 and 8 warps per row and several GPU counts."""
import numpy as np
import pytest

from crow_model import CrowModel
from hg_c_benchmarks_b200 import synth


@pytest.mark.parametrize("n,seg,world,rows,weight,warps",
                         [(1, 32, 1, 1, 1, 4), (31, 32, 1, 2, 1, 4), (33, 32, 1, 2, 3, 4), (65, 64, 1, 2, 1, 4),
                          (200, 32, 1, 1, 1, 4), (200, 64, 1, 3, 2, 4), (333, 96, 1, 2, 5, 4), (300, 8192, 1, 16, 16, 4),
                          (450, 64, 2, 2, 1, 4), (450, 32, 3, 1, 4, 4), (420, 160, 4, 3, 2, 4), (500, 128, 2, 4, 3, 4),
                          (600, 224, 2, 5, 1, 4), (600, 96, 3, 2, 7, 4), (700, 352, 1, 4, 2, 4), (900, 8192, 1, 8, 32, 4),
                          (333, 96, 1, 2, 5, 8), (500, 128, 2, 4, 3, 8), (700, 352, 1, 4, 2, 8), (900, 8192, 2, 8, 32, 8),
                          (641, 160, 1, 3, 1, 8)])
def test_schedule_reproduces_the_in_order_sum(n, seg, world, rows, weight, warps):
    from oracle import port
    pos = synth.uniform_cube(n, seed=n)
    if n >= 64:
        pos[5] = pos[40]                    # duplicate body: d = +0 pairs
        pos[63, 0] = pos[3, 0]
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    m = CrowModel([tuple(map(float, p)) for p in pos], seg, world, rows, weight, warps)
    got = np.array(m.step(), dtype=np.float64).reshape(n, 3)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("n,seg,world,rows,weight,warps,deal", [(450, 64, 2, 1, 4, 4, 4), (600, 96, 3, 2, 7, 4, 6), (700, 128, 4, 1, 16, 4, 3),
                                                                (500, 64, 2, 2, 3, 4, 5), (333, 32, 8, 1, 2, 8, 2)])
def test_rows_handed_out_singly_but_dealt_to_gpus_in_turns(n, seg, world, rows, weight, warps, deal):
    """rows per tile (consecutive tickets) and rows per turn of a GPU are separate knobs"""
    from oracle import port
    pos = synth.uniform_cube(n, seed=n + 1)
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    m = CrowModel([tuple(map(float, p)) for p in pos], seg, world, rows, weight, warps, deal)
    got = np.array(m.step(), dtype=np.float64).reshape(n, 3)
    assert np.array_equal(got, want)
