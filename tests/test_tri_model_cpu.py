"""The second-generation pair-symmetric nbody schedule (csrc/nbody_tri.cuh) checked on the CPU:
tests/tri_model.py mirrors the kernel's index logic (rows, segments, anti-diagonal tile order dealt over
GPUs, minitile row/column passes, the row triangles, tagged hand-offs); its result must be bit-identical
to the oracle's in-order sum (accel(), reference src/nbody/nbody-run.c:17-25) for ragged sizes, several
segment sizes and GPU counts, duplicated bodies and a partial last row."""
import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth
from tri_model import TriModel


@pytest.mark.parametrize("n,seg,world", [(1, 32, 1), (3, 64, 1), (31, 32, 1), (32, 32, 1), (33, 32, 1), (64, 32, 1),
                                         (65, 64, 1), (200, 32, 1), (200, 64, 1), (333, 96, 1), (300, 8192, 1),
                                         (257, 32, 2), (450, 64, 2), (450, 32, 3), (420, 64, 4), (129, 32, 8),
                                         (500, 128, 2), (97, 32, 4)])
def test_schedule_reproduces_the_in_order_sum(n, seg, world):
    from oracle import port
    pos = synth.uniform_cube(n, seed=n)
    if n >= 64:
        pos[5] = pos[40]                    # duplicate body: d = +0 pairs
        pos[63, 0] = pos[3, 0]
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    m = TriModel([tuple(map(float, p)) for p in pos], seg, world)
    got = np.array(m.step(), dtype=np.float64).reshape(n, 3)
    assert np.array_equal(got, want)


def test_plan_is_a_partition_in_dependency_order():
    """every tile (b, k), k <= b, appears on exactly one GPU, after (b, k-1) and (b-1, k)"""
    pos = [tuple(map(float, p)) for p in synth.uniform_cube(700, seed=3)]
    for world in (1, 2, 3, 8):
        m = TriModel(pos, 64, world)
        where = {}
        for g in range(world):
            for t, (b, k) in enumerate(m.plan(g)):
                assert (b, k) not in where and 0 <= k <= b < m.TS and b % world == g
                where[(b, k)] = (b + k, g, t)
        assert len(where) == m.TS * (m.TS + 1) // 2
        for (b, k), (d, g, t) in where.items():
            for dep in ((b, k - 1), (b - 1, k)):
                if dep in where:
                    dd, dg, dt = where[dep]
                    assert dd < d and (dg != g or dt < t)
