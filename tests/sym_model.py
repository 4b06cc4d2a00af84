"""Lock-step Python model of the pair-symmetric nbody schedule (csrc/nbody_sym.cuh).

Test infrastructure: it mirrors, loop for loop, the index logic of `nbody_sym_kernel`
(systolic skew over the 32 lanes, three-chunk ring, open/close of source chunks, the diagonal
block, partial last block) with Python floats (IEEE binary64, one rounding per operator), so
the schedule can be checked on a CPU against the plain in-order sum of the reference
(`accel`, reference src/nbody/nbody-run.c:17-25): every target must receive exactly the same
additions in exactly the same order, hence bit-identical results.

Workers are executed one after the other in ticket order, which is one legal execution of
the device schedule (worker w only ever waits for worker w-1).
"""
import math

G = 9.8      # reference src/nbody/nbody-run.h:12
EPS = 0.005  # reference src/nbody/nbody-run.h:13
LANES = 32
CH = 32     # sources per hand-off chunk


def pair(px, py, pz, xj, yj, zj):
    """d * f of pair_wise_accel (reference src/nbody/nbody-run.c:7-14) for target p, source j."""
    dx, dy, dz = xj - px, yj - py, zj - pz
    s = ((dx * dx + dy * dy) + dz * dz) + EPS
    f = 1.0 / math.sqrt((s * s) * s)
    return dx * f, dy * f, dz * f


def reference_step(pos):
    n = len(pos)
    out = []
    for i in range(n):
        ax = ay = az = 0.0
        for j in range(n):
            vx, vy, vz = pair(*pos[i], *pos[j])
            ax += vx; ay += vy; az += vz
        out.append((G * ax, G * ay, G * az))
    return out


class SymModel:
    def __init__(self, pos, R=2, U=2):
        self.pos, self.n, self.R, self.U = pos, len(pos), R, U
        self.BS = LANES * R
        self.T = (self.n + self.BS - 1) // self.BS
        npad = self.T * self.BS
        self.x = [p[0] for p in pos] + [0.0] * (npad - self.n)
        self.y = [p[1] for p in pos] + [0.0] * (npad - self.n)
        self.z = [p[2] for p in pos] + [0.0] * (npad - self.n)
        self.acc = [[None] * npad for _ in range(3)]   # global accumulator chain (SoA)
        self.nchunks = (self.n + CH - 1) // CH
        self.done = [-1] * self.nchunks                # done[c] = last worker that closed chunk c
        self.fast_steps = self.slow_steps = 0

    # -- chunk traffic (one chunk = 32 sources, the hand-off unit between workers) -------------
    def open(self, w, c, ring):
        base = (c % 3) * CH
        lower = c < w * self.R
        if lower:
            assert self.done[c] == w - 1, (w, c, self.done[c])
        for idx in range(CH):
            g = c * CH + idx
            ring[0][base + idx], ring[1][base + idx], ring[2][base + idx] = self.x[g], self.y[g], self.z[g]
            for k in range(3):
                ring[3 + k][base + idx] = self.acc[k][g] if lower else None

    def close(self, w, c, ring):
        base = (c % 3) * CH
        for idx in range(CH):
            for k in range(3):
                self.acc[k][c * CH + idx] = ring[3 + k][base + idx]
        self.done[c] = w

    # -- one worker = one warp ----------------------------------------------------------
    def worker(self, w):
        R, U, BS, n = self.R, self.U, self.BS, self.n
        RING = 3 * CH
        i0 = w * BS
        jend = min(i0 + BS, n)
        ring = [[None] * RING for _ in range(6)]
        tgt = [[i0 + R * l + r for r in range(R)] for l in range(LANES)]
        acc = [[[0.0, 0.0, 0.0] for _ in range(R)] for _ in range(LANES)]
        e = [(RING - l) % RING for l in range(LANES)]

        def fast_step(s):
            for l in range(LANES):
                j = s - l
                assert 0 <= j < i0 and e[l] == j % RING
                u = [ring[3][e[l]], ring[4][e[l]], ring[5][e[l]]]
                for r in range(R):
                    i = tgt[l][r]
                    v = pair(self.x[i], self.y[i], self.z[i], ring[0][e[l]], ring[1][e[l]], ring[2][e[l]])
                    for c in range(3):
                        acc[l][r][c] = acc[l][r][c] + v[c]
                        if i < n:
                            u[c] = u[c] - v[c]
                ring[3][e[l]], ring[4][e[l]], ring[5][e[l]] = u
                e[l] = 0 if e[l] + 1 == RING else e[l] + 1

        def slow_step(s):
            for l in range(LANES):
                j = s - l
                if 0 <= j < jend:
                    assert e[l] == j % RING
                    have_u = j < tgt[l][0]
                    u = [ring[3][e[l]], ring[4][e[l]], ring[5][e[l]]] if have_u else [None] * 3
                    for r in range(R):
                        i = tgt[l][r]
                        if i < n:
                            if j == i:
                                u, have_u = list(acc[l][r]), True
                            elif j < i:
                                v = pair(self.x[i], self.y[i], self.z[i], ring[0][e[l]], ring[1][e[l]], ring[2][e[l]])
                                for c in range(3):
                                    acc[l][r][c] = acc[l][r][c] + v[c]
                                    u[c] = u[c] - v[c]
                    if have_u:
                        ring[3][e[l]], ring[4][e[l]], ring[5][e[l]] = u
                e[l] = 0 if e[l] + 1 == RING else e[l] + 1

        s, s_end, last_chunk = 0, jend - 1 + (LANES - 1), (jend - 1) // CH
        while s <= s_end:
            if s % CH == 0 and s // CH <= last_chunk:
                c = s // CH
                if c >= 2:
                    self.close(w, c - 2, ring)
                self.open(w, c, ring)
            lim = min(i0, (s // CH + 1) * CH)
            if s >= LANES - 1 and s + U <= lim:
                while s + U <= lim:
                    for u_ in range(U):
                        fast_step(s + u_)
                    self.fast_steps += U
                    s += U
            else:                                   # edge_group: U predicated steps, s stays a multiple of U
                assert s % U == 0 and (s % CH) + U <= CH
                for u_ in range(U):
                    slow_step(s + u_)
                self.slow_steps += U
                s += U
        if last_chunk >= 1:
            self.close(w, last_chunk - 1, ring)
        self.close(w, last_chunk, ring)

    def step(self):
        for w in range(self.T):
            self.worker(w)
        assert self.done == [self.T - 1] * self.nchunks
        return [(G * self.acc[0][i], G * self.acc[1][i], G * self.acc[2][i]) for i in range(self.n)]
