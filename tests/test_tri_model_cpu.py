"""The second-generation pair-symmetric nbody schedule (csrc/nbody_tri.cuh) checked on the CPU:
tests/tri_model.py mirrors the kernel's index logic (rows, segments, anti-diagonal tile order dealt over
GPUs, minitile row/column passes, the row triangles, tagged hand-offs); its result must be bit-identical
to the oracle's in-order sum (accel(), reference src/nbody/nbody-run.c:17-25) for ragged sizes, several
segment sizes and GPU counts, duplicated bodies and a partial last row."""
import numpy as np
import pytest

from hg_c_benchmarks_b200 import synth
from tri_model import TriModel


@pytest.mark.parametrize("n,seg,world,rows,weight",
                         [(1, 32, 1, 1, 1), (3, 64, 1, 2, 1), (31, 32, 1, 16, 16), (32, 32, 1, 1, 2), (33, 32, 1, 2, 3),
                          (64, 32, 1, 1, 1), (65, 64, 1, 2, 1), (200, 32, 1, 1, 1), (200, 64, 1, 3, 2), (333, 96, 1, 2, 5),
                          (300, 8192, 1, 16, 16), (257, 32, 2, 1, 1), (450, 64, 2, 2, 1), (450, 32, 3, 1, 4),
                          (420, 64, 4, 3, 2), (129, 32, 8, 1, 1), (500, 128, 2, 4, 3), (97, 32, 4, 2, 2),
                          (600, 64, 2, 5, 1), (600, 96, 3, 2, 7)])
def test_schedule_reproduces_the_in_order_sum(n, seg, world, rows, weight):
    from oracle import port
    pos = synth.uniform_cube(n, seed=n)
    if n >= 64:
        pos[5] = pos[40]                    # duplicate body: d = +0 pairs
        pos[63, 0] = pos[3, 0]
    want = np.zeros_like(pos)
    port.nbody_step_range(pos, want, 0, n)
    m = TriModel([tuple(map(float, p)) for p in pos], seg, world, rows, weight)
    got = np.array(m.step(), dtype=np.float64).reshape(n, 3)
    assert np.array_equal(got, want)


def test_plan_is_a_partition_in_dependency_order():
    """every item (I, k) runs on exactly one GPU, after (I, k-1) and after the item of row I-1 that
    publishes the chain words it reads: (I-1, k), or (I-1, k-1) where row I-1 ends before segment k"""
    pos = [tuple(map(float, p)) for p in synth.uniform_cube(900, seed=3)]
    for world, seg, rows, weight in ((1, 64, 2, 1), (2, 64, 1, 3), (3, 96, 2, 2), (8, 64, 3, 1), (2, 128, 5, 4)):
        m = TriModel(pos, seg, world, rows, weight)
        when = {}                                   # item -> (global key, shard, position in the shard's ticket stream)
        for g in range(world):
            t = 0
            for (b, k) in m.plan(g):
                assert b % world == g
                for r in range(rows):
                    I = b * rows + r
                    if I < m.T and k <= 32 * I // seg:
                        assert (I, k) not in when
                        when[(I, k)] = ((b + weight * k) * 4096 + k, g, t)
                    t += 1
        assert len(when) == sum(32 * I // seg + 1 for I in range(m.T))
        for (I, k), (key, g, t) in when.items():
            deps = [(I, k - 1)] if k > 0 else []
            if I > 0 and 32 * I > k * seg:          # the item has lower sources in segment k
                deps.append((I - 1, k) if (I - 1, k) in when else (I - 1, k - 1))
            for dep in deps:
                dkey, dg, dt = when[dep]
                assert dkey < key or (dkey == key and dg == g and dt < t), ((I, k), dep)
