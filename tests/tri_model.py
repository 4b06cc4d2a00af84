"""CPU model of the second-generation pair-symmetric nbody schedule (csrc/nbody_tri.cuh).

Test infrastructure: it mirrors the index logic of `nbody_tri_kernel` -- rows of 32 targets, segments,
tiles of `rows` rows in key order (tile-row + weight * segment) dealt over `world` GPUs, 32 x 8 minitiles with a row pass (targets add
their 8 sources in order) and a column pass (each source's chain subtracts the 32 targets in order), the
row's own triangle with chain seeds, tagged hand-off words between rows and between the items of a row,
the final row's broadcast -- with Python floats (IEEE binary64, one rounding per operator).  Items are
executed one after the other in global ticket order, which is one legal execution of the device
schedule; every tag the kernel would poll for is asserted to be there already, which is exactly the
"only ever waits for an earlier ticket" property.  Results must equal the plain in-order sum of the
reference (`accel`, reference src/nbody/nbody-run.c:17-25) bit for bit.
"""
from sym_model import G, pair

LANES = 32
MT = 8


class TriModel:
    def __init__(self, pos, seg=64, world=1, rows=2, weight=3, deal=None):
        assert seg % 32 == 0 and rows >= 1 and weight >= 1
        self.deal = rows if deal is None else max(rows, deal // rows * rows)   # rows per turn of a GPU, a multiple of the rows per tile
        self.pos, self.n, self.seg, self.world, self.weight = pos, len(pos), seg, world, weight
        self.T = (self.n + 31) // 32
        self.tr = rows                                    # rows per tile
        self.TS = (self.T + self.tr - 1) // self.tr       # tile-rows
        npad = self.T * 32
        self.x = [p[0] for p in pos] + [0.0] * (npad - self.n)
        self.y = [p[1] for p in pos] + [0.0] * (npad - self.n)
        self.z = [p[2] for p in pos] + [0.0] * (npad - self.n)
        # per GPU: chain words ll[j] = (tag, [x, y, z]); row accumulators racc[I] = (tag, [[x,y,z]] * 32)
        self.ll = [[(0, None)] * npad for _ in range(world)]
        self.racc = [[(0, None)] * self.T for _ in range(world)]
        self.items = 0

    def plan(self, index):
        """tiles of shard `index` in ticket order (hgb_nbody.cu tri_plan): key b + weight * k, then k"""
        tiles = []
        for b in range(self.TS):
            if (b * self.tr // self.deal) % self.world != index:
                continue
            last_row = min(self.T - 1, b * self.tr + self.tr - 1)
            for k in range(32 * last_row // self.seg + 1):
                tiles.append(((b + self.weight * k) * 4096 + k, b, k))
        return [(b, k) for _, b, k in sorted(tiles)]

    def item(self, dev, b, k, r):
        I = b * self.tr + r
        if I >= self.T:
            return
        n, i0 = self.n, 32 * I
        K = i0 // self.seg                                # the row's last item
        if k > K:
            return                                        # this row ends before segment k
        self.items += 1
        last_item = k == K
        j_begin, j_end = k * self.seg, (i0 if last_item else (k + 1) * self.seg)
        cnt = min(32, n - i0)
        final_row = I == self.T - 1
        want_tag, my_tag = I, I + 1
        # where row I+1 reads: same GPU, or the next super-row's GPU after the last row of a super-row
        pub = [(dev + 1) % self.world] if I % self.deal == self.deal - 1 else [dev]
        if final_row:
            pub = list(range(self.world))
        if k > 0:
            tag, acc = self.racc[dev][I]
            assert tag == k, ("row accumulators not there yet", I, k, tag)
            acc = [list(a) for a in acc]
        else:
            acc = [[0.0, 0.0, 0.0] for _ in range(LANES)]
        tgt = [(self.x[i0 + l], self.y[i0 + l], self.z[i0 + l]) for l in range(LANES)]

        assert (j_end - j_begin) % 32 == 0
        for j0 in range(j_begin, j_end, MT):              # lower minitiles, all 32 x 8 pairs
            v = [[pair(*tgt[l], self.x[j0 + u], self.y[j0 + u], self.z[j0 + u]) for u in range(MT)] for l in range(LANES)]
            for l in range(LANES):                        # row pass
                for u in range(MT):
                    for c in range(3):
                        acc[l][c] = acc[l][c] + v[l][u][c]
            for u in range(MT):                           # column pass
                tag, chain = self.ll[dev][j0 + u]
                assert tag == want_tag, ("chain word not there yet", I, k, j0 + u, tag)
                chain = list(chain)
                for l in range(cnt):
                    for c in range(3):
                        chain[c] = chain[c] - v[l][u][c]
                for p in pub:
                    self.ll[p][j0 + u] = (my_tag, chain)
        if not last_item:
            self.racc[dev][I] = (k + 1, acc)
            return
        for mt in range(4):                               # the row's own triangle
            v = [[pair(*tgt[l], *tgt[mt * MT + u]) for u in range(MT)] for l in range(LANES)]
            for l in range(LANES):
                for u in range(MT):
                    if mt * MT + u < l:
                        for c in range(3):
                            acc[l][c] = acc[l][c] + v[l][u][c]
            for u in range(MT):
                chain = list(acc[mt * MT + u])            # seed: the complete lower sum of that body
                for l in range(mt * MT + u + 1, cnt):
                    for c in range(3):
                        chain[c] = chain[c] - v[l][u][c]
                for p in pub:
                    self.ll[p][i0 + mt * MT + u] = (my_tag, chain)

    def step(self):
        plans = [self.plan(g) for g in range(self.world)]
        # global key order: its restriction to a shard is that shard's ticket order
        order = sorted((((b + self.weight * k) * 4096 + k, b, k, g) for g in range(self.world) for (b, k) in plans[g]))
        seen = set()
        for _, b, k, g in order:
            assert (b * self.tr // self.deal) % self.world == g and (b, k) not in seen
            seen.add((b, k))
            for r in range(self.tr):
                self.item(g, b, k, r)
        out = []
        for g in range(self.world):                       # every GPU holds the finished sums
            res = []
            for i in range(self.n):
                tag, a = self.ll[g][i]
                assert tag == self.T, (g, i, tag)
                res.append((G * a[0], G * a[1], G * a[2]))
            out.append(res)
        for g in range(1, self.world):
            assert out[g] == out[0]
        return out[0]
