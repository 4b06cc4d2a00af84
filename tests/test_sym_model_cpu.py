"""The pair-symmetric nbody schedule (csrc/nbody_sym.cuh) checked on the CPU: tests/sym_model.py
mirrors the kernel's index logic lane by lane; its result must be bit-identical to the oracle's
in-order sum (accel(), reference src/nbody/nbody-run.c:17-25) for ragged sizes, every (R, U)
the library instantiates, duplicated bodies and a partial last block."""
import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth
from sym_model import SymModel, reference_step


@pytest.mark.parametrize("n,R,U", [(1, 2, 2), (3, 2, 2), (63, 2, 2), (64, 2, 2), (65, 2, 2), (200, 2, 2),
                                   (333, 2, 2), (300, 1, 2), (300, 1, 4), (257, 2, 1), (450, 2, 4),
                                   (300, 4, 1), (420, 4, 2), (129, 4, 2)])
def test_schedule_reproduces_the_in_order_sum(n, R, U):
    from oracle import port
    pos = synth.uniform_cube(n, seed=n)
    if n >= 64:
        pos[5] = pos[40]                    # duplicate body: d = +0 pairs
        pos[63, 0] = pos[3, 0]
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    m = SymModel([tuple(map(float, p)) for p in pos], R, U)
    got = np.array(m.step(), dtype=np.float64).reshape(n, 3)
    assert np.array_equal(got, want)
    assert np.array_equal(np.array(reference_step([tuple(map(float, p)) for p in pos])).reshape(n, 3), want)
    if n > 32 * R + 40:
        assert m.fast_steps > 0             # the predicate-free path was exercised


def test_every_pair_is_evaluated_once():
    """work count: n(n-1)/2 pair evaluations, vs n^2 interactions of the plain schedule"""
    import sym_model
    calls = [0]
    orig = sym_model.pair

    def counting(*a):
        calls[0] += 1
        return orig(*a)
    sym_model.pair = counting
    try:
        n = 200
        pos = [tuple(map(float, p)) for p in synth.uniform_cube(n, seed=1)]
        SymModel(pos, 2, 2).step()
    finally:
        sym_model.pair = orig
    # padding targets of the last block evaluate (and discard) pairs on the predicate-free path
    assert n * (n - 1) // 2 <= calls[0] <= n * (n - 1) // 2 + 64 * n
