"""Lock-step Python model of the segmented first-generation schedule (csrc/nbody_seg.cuh).

Test infrastructure.  tests/sym_model.py mirrors the systolic warp of `nbody_sym_kernel`; this model keeps its
lane / ring / chunk logic and adds what `nbody_seg_kernel` adds: a block's row is cut into ITEMS (block w against
segment k of `seg` sources), an item's stream starts at j_begin = k seg and stops at j_stop (the segment's end, or the
end of the block itself in the row's last item), the skew fills and drains once per item, the block's row sums travel
from item (w, k-1) to (w, k) as tagged words, chunks are opened / closed per item (closed two chunks behind, the last
but one also as soon as every lane has left it), and the items are run in the ticket order of the host plan (tiles of
`rows` blocks in key order tile + weight * segment, dealt over `world` GPUs in turns of `deal` blocks).  The software
pipeline of the device kernel (forces of group g+1 evaluated while the ordered steps of group g are applied) moves no
addition: the steps of a group still run in order after its forces and before the next group's steps, which is the
order modelled here.  Every hand-off the kernel would poll for is asserted to be there already ("only ever waits for
an earlier ticket").  Python floats (binary64, one rounding per operator); the result must equal the in-order sum of
accel() (reference src/nbody/nbody-run.c:17-25) bit for bit.
"""
from sym_model import CH, G, LANES, pair


class SegModel:
    def __init__(self, pos, seg=64, rows=2, weight=3, world=1, deal=None, R=2, U=4):
        assert seg % CH == 0 and rows >= 1 and weight >= 1
        self.pos, self.n, self.R, self.U, self.seg, self.tr, self.weight, self.world = pos, len(pos), R, U, seg, rows, weight, world
        self.deal = rows if deal is None else max(rows, deal // rows * rows)
        self.BS = LANES * R
        self.T = (self.n + self.BS - 1) // self.BS
        npad = self.T * self.BS
        self.x = [p[0] for p in pos] + [0.0] * (npad - self.n)
        self.y = [p[1] for p in pos] + [0.0] * (npad - self.n)
        self.z = [p[2] for p in pos] + [0.0] * (npad - self.n)
        # per GPU: chain words per source (tag = producing block + 1, value), row sums per block (tag = next item, values)
        self.ll = [[(0, None)] * npad for _ in range(world)]
        self.racc = [[(0, None)] * self.T for _ in range(world)]
        self.items = 0

    def owner(self, w):
        return (w // self.deal) % self.world

    def plan(self, index):
        """tiles of shard `index` in ticket order (hgb_nbody.cu seg_plan / tri_plan with 64-target rows)"""
        TB = (self.T + self.tr - 1) // self.tr
        tiles = []
        for b in range(TB):
            if (b * self.tr // self.deal) % self.world != index:
                continue
            last_row = min(self.T - 1, b * self.tr + self.tr - 1)
            for k in range(self.BS * last_row // self.seg + 1):
                tiles.append(((b + self.weight * k) * 4096 + k, b, k))
        return [(b, k) for _, b, k in sorted(tiles)]

    # -- one item = one warp's pass over one segment ------------------------------------------------
    def item(self, dev, w, k):
        R, U, BS, n, seg = self.R, self.U, self.BS, self.n, self.seg
        RING = 3 * CH
        i0 = w * BS
        K = i0 // seg
        if w >= self.T or k > K:
            return
        self.items += 1
        last_item = k == K
        jend = min(i0 + BS, n)
        j_begin, j_stop = k * seg, (jend if last_item else (k + 1) * seg)
        want_tag, my_tag = w, w + 1
        ll = self.ll[dev]
        if w == self.T - 1:
            pubs = self.ll                                         # model shortcut for the finish kernels' peer reads
        elif w % self.deal == self.deal - 1:
            pubs = [self.ll[(dev + 1) % self.world]]
        else:
            pubs = [ll]
        tgt = [[i0 + R * l + r for r in range(R)] for l in range(LANES)]
        if k > 0:
            tag, vals = self.racc[dev][w]
            assert tag == k, ("row sums not there yet", w, k, tag)
            acc = [[list(v) for v in lane] for lane in vals]
        else:
            acc = [[[0.0, 0.0, 0.0] for _ in range(R)] for _ in range(LANES)]
        ring = [[None] * RING for _ in range(6)]
        e = [((j_begin - l) % RING + RING) % RING for l in range(LANES)]

        def open_chunk(c):
            base = (c % 3) * CH
            lower = c < w * R
            for idx in range(CH):
                g = c * CH + idx
                ring[0][base + idx], ring[1][base + idx], ring[2][base + idx] = self.x[g], self.y[g], self.z[g]
                if lower:
                    tag, val = ll[g]
                    assert tag == want_tag, ("chain word not there yet", w, k, g, tag)
                    for comp in range(3):
                        ring[3 + comp][base + idx] = val[comp]
                else:
                    for comp in range(3):
                        ring[3 + comp][base + idx] = None

        def close_chunk(c):
            base = (c % 3) * CH
            for idx in range(CH):
                val = [ring[3 + comp][base + idx] for comp in range(3)]
                for p in pubs:
                    p[c * CH + idx] = (my_tag, val)

        def step(s, fast):
            for l in range(LANES):
                j = s - l
                act = j_begin <= j < j_stop
                if fast:
                    assert act and j < i0
                if act:
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
                                for comp in range(3):
                                    acc[l][r][comp] = acc[l][r][comp] + v[comp]
                                    u[comp] = u[comp] - v[comp]
                    if have_u:
                        ring[3][e[l]], ring[4][e[l]], ring[5][e[l]] = u
                e[l] = 0 if e[l] + 1 == RING else e[l] + 1

        s, early = j_begin, False
        s_end, first_chunk, last_chunk = j_stop - 1 + (LANES - 1), j_begin // CH, (j_stop - 1) // CH
        fast_from, fast_to = j_begin + LANES - 1, min(i0, j_stop)
        while s <= s_end:
            if s % CH == 0 and s // CH <= last_chunk:
                c = s // CH
                if c - 2 >= first_chunk:
                    close_chunk(c - 2)
                open_chunk(c)
            elif s == (last_chunk + 1) * CH and last_chunk - 1 >= first_chunk and not early:
                close_chunk(last_chunk - 1)
                early = True
            lim = min(fast_to, (s // CH + 1) * CH)
            if s >= fast_from and s + U <= lim:
                while s + U <= lim:                 # the pipelined run of the device kernel: same steps, same order
                    for u_ in range(U):
                        step(s + u_, True)
                    s += U
            else:
                assert s % U == 0
                for u_ in range(U):
                    step(s + u_, False)
                s += U
        if last_chunk - 1 >= first_chunk and not early:
            close_chunk(last_chunk - 1)
        close_chunk(last_chunk)
        if not last_item:
            self.racc[dev][w] = (k + 1, acc)

    def step(self):
        plans = [self.plan(g) for g in range(self.world)]
        order = sorted((((b + self.weight * k) * 4096 + k, b, k, g) for g in range(self.world) for (b, k) in plans[g]))
        seen = set()
        for _, b, k, g in order:
            assert (b * self.tr // self.deal) % self.world == g and (b, k) not in seen
            seen.add((b, k))
            for r in range(self.tr):
                self.item(g, b * self.tr + r, k)
        out = []
        for g in range(self.world):
            res = []
            for i in range(self.n):
                tag, a = self.ll[g][i]
                assert tag == self.T, (g, i, tag)
                res.append((G * a[0], G * a[1], G * a[2]))
            out.append(res)
        for g in range(1, self.world):
            assert out[g] == out[0]
        return out[0]
