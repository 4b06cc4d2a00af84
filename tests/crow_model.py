"""CPU model of the third-generation pair-symmetric nbody schedule (csrc/nbody_crow.cuh): a CTA of W warps
per row.

Test infrastructure.  Where tests/tri_model.py models the schedule item by item, this one goes down to the
index arithmetic of `nbody_crow_kernel`'s inner loops, because that is what changed: minitile m of an item
belongs to warp m mod W; every warp keeps its own two-slot source ring (chunks of 4 of ITS minitiles, loaded
lane-wise with the kernel's `src_of` gather), its own two transposed buffers, its two chain-word registers
(`we`/`wo`, refilled two minitiles ahead, addressed by pointer arithmetic in 16-byte word units exactly as in
the kernel: word = chunk * 96 + component * 32 + source % 32), the peeled first minitile, the (odd, even) pair
loop, the optional trailing odd minitile and the final column pass; the row's sums travel from warp to warp
as tokens tagged with the number of minitiles absorbed.  Warps are Python generators that yield while a
token or chain word is missing, so any interleaving the device could produce that respects the hand-offs is
emulated by a round-robin scheduler, and a hand-off that can never arrive shows up as a stall (assert).
Arithmetic is Python floats (binary64, one rounding per operator); the result must equal the in-order sum
of accel() (reference src/nbody/nbody-run.c:17-25) bit for bit.
"""
from sym_model import G, pair
from tri_model import TriModel

PITCH = 34


class CrowModel(TriModel):
    def __init__(self, pos, seg=64, world=1, rows=2, weight=3, warps=4, deal=None):
        super().__init__(pos, seg, world, rows, weight, deal)
        assert warps % 4 == 0
        self.W = warps
        self.npad = self.T * 32
        nwords = 3 * self.npad
        # flat tagged-word buffers, indexed like the device buffers
        self.llw = [[(0, 0.0)] * nwords for _ in range(world)]
        self.raccw = [[(0, 0.0)] * nwords for _ in range(world)]

    # ---- one warp of one item, as a generator --------------------------------------------------
    def warp(self, dev, q, I, k, r, slots):
        W, seg, n = self.W, self.seg, self.n
        WS = 24 * W
        ll, racc = self.llw[dev], self.raccw[dev]
        i0 = 32 * I
        K = i0 // seg
        last_item = k == K
        j_begin, j_end = k * seg, (i0 if last_item else (k + 1) * seg)
        cnt = min(32, n - i0)
        final_row = I == self.T - 1
        want_tag, my_tag = I, I + 1
        if final_row:
            pubs = [self.llw[g] for g in range(self.world)]    # model shortcut for the finish kernel's peer reads
        elif I % self.deal == self.deal - 1:
            pubs = [self.llw[(dev + 1) % self.world]]
        else:
            pubs = [ll]
        M = (j_end - j_begin) >> 3
        C = (M - q + W - 1) // W if M > q else 0
        px = [self.x[i0 + l] for l in range(32)]
        py = [self.y[i0 + l] for l in range(32)]
        pz = [self.z[i0 + l] for l in range(32)]
        acc = [[0.0, 0.0, 0.0] for _ in range(32)]
        state = {"acc": acc}

        def take_token(tok):
            if tok == 0:
                for l in range(32):
                    state["acc"][l] = [0.0, 0.0, 0.0]
                if k > 0:
                    for c in range(3):
                        for l in range(32):
                            while racc[I * 96 + c * 32 + l][0] != k:
                                yield
                            state["acc"][l][c] = racc[I * 96 + c * 32 + l][1]
                return
            prev = (q + W - 1) % W
            for c in range(3):
                for l in range(32):
                    while slots[prev][c][l][0] != tok:
                        yield
                    state["acc"][l][c] = slots[prev][c][l][1]

        def give_token(tok):
            for c in range(3):
                for l in range(32):
                    slots[q][c][l] = (tok, state["acc"][l][c])

        def publish(widx, lanes_chain):
            for L in range(24):
                for p in pubs:
                    p[widx(L)] = (my_tag, lanes_chain[L])

        if C > 0:
            nch = (C + 3) >> 2
            ring = [[0.0] * 64 for _ in range(3)]               # [comp][slot * 32 + lane]
            tb = [[[0.0] * PITCH for _ in range(24)] for _ in range(2)]

            def src_of(ch, lane):
                return min(j_begin + 8 * (q + W * (4 * ch + (lane >> 3))) + (lane & 7), j_end - 1)

            def word0(L):                                       # this column lane's word of minitile 0 (16-byte units)
                cc, cu = L >> 3, L & 7
                return ((j_begin >> 5) + (q >> 2)) * 96 + cc * 32 + (q & 3) * 8 + cu

            for l in range(32):
                g = src_of(0, l)
                ring[0][l], ring[1][l], ring[2][l] = self.x[g], self.y[g], self.z[g]
            f = [None] * 32
            if nch > 1:
                f = [(self.x[src_of(1, l)], self.y[src_of(1, l)], self.z[src_of(1, l)]) for l in range(32)]
            wp = 0                                              # offset from word0, in word units
            # registers: we/wo hold, per column lane, the word INDEX they were loaded from (the load is volatile
            # and re-polled until the tag shows, so modelling the address is what matters)
            we = wp
            wo = wp + WS if C > 1 else None
            src = {"x": None}

            def fetch_sources(i):
                base = (i & 7) * 8
                src["x"] = [(ring[0][base + u], ring[1][base + u], ring[2][base + u]) for u in range(8)]

            def next_chunk(ch):
                nonlocal f
                if ch + 1 < nch:
                    o = ((ch + 1) & 1) * 32
                    for l in range(32):
                        ring[0][o + l], ring[1][o + l], ring[2][o + l] = f[l]
                    if ch + 2 < nch:
                        f = [(self.x[src_of(ch + 2, l)], self.y[src_of(ch + 2, l)], self.z[src_of(ch + 2, l)]) for l in range(32)]

            def check_sources(i):
                """the ring must hold exactly the sources of global minitile q + W i"""
                m = q + W * i
                for u in range(8):
                    j = j_begin + 8 * m + u
                    assert src["x"][u] == (self.x[j], self.y[j], self.z[j]), ("ring holds the wrong sources", I, k, q, i, u)

            def forces(i):
                check_sources(i)
                return [[pair(px[l], py[l], pz[l], *src["x"][u]) for u in range(8)] for l in range(32)]

            def wait_words(at):
                """chain words of the minitile whose word offset is `at`: per column lane"""
                out = []
                for L in range(24):
                    idx = word0(L) + at
                    while ll[idx][0] != want_tag:
                        yield None
                    out.append(ll[idx][1])
                yield out

            def column(chain, buf):
                for L in range(24):
                    for t in range(cnt):
                        chain[L] = chain[L] - tb[buf][L][t]
                return chain

            def row_pass(i, buf, v):
                m = q + W * i
                yield from take_token(m)
                for l in range(32):
                    for u in range(8):
                        for c in range(3):
                            state["acc"][l][c] = state["acc"][l][c] + v[l][u][c]
                give_token(m + 1)
                for l in range(32):
                    for u in range(8):
                        for c in range(3):
                            tb[buf][c * 8 + u][l] = v[l][u][c]

            def block(i, reg_at, at, colbuf, minebuf, more):
                assert reg_at == at, ("word register holds the wrong minitile", I, k, q, i, reg_at, at)
                chain = None
                for res in wait_words(at):
                    if res is None:
                        yield
                    else:
                        chain = res
                v = forces(i)
                chain = column(chain, colbuf)
                if more:
                    fetch_sources(i + 1)
                yield from row_pass(i, minebuf, v)
                publish(lambda L: word0(L) + at, chain)

            fetch_sources(0)
            next_chunk(0)
            v = forces(0)
            if C > 1:
                fetch_sources(1)
            yield from row_pass(0, 0, v)
            i = 1
            while i + 1 < C:
                yield from block(i, we, wp, 0, 1, True)
                we = wp + 2 * WS
                if (i & 3) == 3:
                    next_chunk((i + 1) >> 2)
                more = i + 2 < C
                yield from block(i + 1, wo, wp + WS, 1, 0, more)
                if more:
                    wo = wp + 3 * WS
                wp += 2 * WS
                i += 2
            last_buf = 0
            if i < C:
                yield from block(i, we, wp, 0, 1, False)
                wp += WS
                last_buf = 1
            reg = we if last_buf == 0 else wo
            assert reg == wp, ("final word register holds the wrong minitile", I, k, q, reg, wp)
            # the word offset of this warp's last minitile must be the one of global minitile q + W (C-1)
            m_last = q + W * (C - 1)
            for L in range(24):
                assert word0(L) + wp == ((j_begin >> 5) + (m_last >> 2)) * 96 + (L >> 3) * 32 + (m_last & 3) * 8 + (L & 7)
            chain = None
            for res in wait_words(wp):
                if res is None:
                    yield
                else:
                    chain = res
            chain = column(chain, last_buf)
            publish(lambda L: word0(L) + wp, chain)

        if q == M % W:
            yield from take_token(M)
            acc = state["acc"]
            if not last_item:
                for c in range(3):
                    for l in range(32):
                        racc[I * 96 + c * 32 + l] = (k + 1, acc[l][c])
                return
            tgt = list(zip(px, py, pz))
            for mt in range(4):
                v = [[pair(*tgt[l], *tgt[mt * 8 + u]) for u in range(8)] for l in range(32)]
                for l in range(32):
                    for u in range(8):
                        if mt * 8 + u < l:
                            for c in range(3):
                                acc[l][c] = acc[l][c] + v[l][u][c]
                for L in range(24):
                    cc, cu = L >> 3, L & 7
                    chain = acc[mt * 8 + cu][cc]
                    for t in range(mt * 8 + cu + 1, cnt):
                        chain = chain - v[t][cu][cc]
                    for p in pubs:
                        p[I * 96 + cc * 32 + mt * 8 + cu] = (my_tag, chain)

    def item(self, dev, b, k, r):
        I = b * self.tr + r
        if I >= self.T or k > 32 * I // self.seg:
            return
        self.items += 1
        slots = [[[(0, 0.0)] * 32 for _ in range(3)] for _ in range(self.W)]
        warps = [self.warp(dev, q, I, k, r, slots) for q in range(self.W)]
        alive = list(range(self.W))
        idle_rounds = 0
        while alive:
            progressed = False
            for q in list(alive):
                try:
                    # run the warp until it has to wait (one `yield` = one failed poll); a warp that keeps
                    # polling without anyone else moving is a lost hand-off
                    next(warps[q])
                except StopIteration:
                    alive.remove(q)
                    progressed = True
            # each next() either finishes or stops at a failed poll; progress = somebody finished or some slot changed
            snapshot = repr([[s[0][0][0] for s in slots]])
            if not progressed and getattr(self, "_snap", None) == snapshot:
                idle_rounds += 1
                assert idle_rounds < 4, ("warps stalled: a hand-off never arrives", I, k, alive)
            else:
                idle_rounds = 0
            self._snap = snapshot

    def step(self):
        plans = [self.plan(g) for g in range(self.world)]
        order = sorted((((b + self.weight * k) * 4096 + k, b, k, g) for g in range(self.world) for (b, k) in plans[g]))
        for _, b, k, g in order:
            for r in range(self.tr):
                self.item(g, b, k, r)
        out = []
        for g in range(self.world):
            res = []
            for i in range(self.n):
                w = [self.llw[g][(i >> 5) * 96 + c * 32 + (i & 31)] for c in range(3)]
                assert all(t == self.T for t, _ in w), (g, i, w)
                res.append((G * w[0][1], G * w[1][1], G * w[2][1]))
            out.append(res)
        for g in range(1, self.world):
            assert out[g] == out[0]
        return out[0]
