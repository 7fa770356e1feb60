"""Happens-before model of the look-ahead schedule (race detection without a GPU).

The engine runs up to five kernels of different steps at the same time on three streams per GPU,
ordered only by CUDA events, by flags in device memory and by the cross-GPU min-loc exchange.  This
module rebuilds that partial order for a small problem and checks that every pair of conflicting
accesses (same buffer tile, at least one write) is ordered:

* nodes   - start / end of every kernel launch and of every work item inside it, event records and
            waits, flag publications;
* edges   - stream order, event record -> wait, "kernel start -> its items -> kernel end", flag
            publication -> waiting item (exactly the waits of csrc/step_fused.cu), K1's min-loc
            (a rank's K1(t) ends after every rank's K1(t) has started);
* items   - taken from the REAL work-list code (csrc/step_plan.h through jb_plan_enumerate), so the
            phase contents and their order are the kernel's own;
* accesses- per item, at tile granularity: W[slot][tile], pivot-row buffer [parity][tile] on every
            rank, L panel [step % 3], candidate inverses [parity], stepinfo[step].

The host-side launch sequence below mirrors jb_engine_invert (csrc/engine.cu), including its
switches (sched: 2 = phase A beside B2, 1 = in line; b2_overlap; k1_gated).  A hazard reported here
is a possible race on the GPU whatever the timing; no report means the model found every conflicting
pair ordered.  Pivot sequences may leave the diagonal: the block-row swap (peer stores into the pivot
owner's swap buffer, swap flags, update of the swapped-in row from that buffer) is part of the model.
"""
from __future__ import annotations

import ctypes as C
from collections import defaultdict

from mpi_jordan_crazy_acceleration_b200 import _native

SCALE, ROWT, UPD, UPD_SW = 0, 1, 2, 3
PH_A, PH_S, PH_B1, PH_B2 = 1, 4, 5, 6
BN = 128


class Graph:
    def __init__(self):
        self.names: list[str] = []
        self.preds: list[list[int]] = []

    def node(self, name: str, *preds: int) -> int:
        self.names.append(name)
        self.preds.append([p for p in preds if p is not None])
        return len(self.names) - 1

    def edge(self, a: int, b: int) -> None:
        if a is not None:
            self.preds[b].append(a)

    def closure(self) -> list[int]:
        """reach[v] = bitset of nodes that happen before v (nodes are created in an order that is
        NOT topological - flags of other ranks - so iterate to a fixed point)."""
        n = len(self.names)
        reach = [0] * n
        changed = True
        while changed:
            changed = False
            for v in range(n):
                r = reach[v]
                for u in self.preds[v]:
                    r |= reach[u] | (1 << u)
                if r != reach[v]:
                    reach[v] = r
                    changed = True
        return reach


class Stream:
    """Work on a stream is serial: every op depends on the previous one."""

    def __init__(self, g: Graph, name: str):
        self.g, self.name, self.tail = g, name, None

    def op(self, name: str, *preds: int) -> int:
        v = self.g.node(f"{self.name}:{name}", self.tail, *preds)
        self.tail = v
        return v


def _enumerate(m, p, rank, rloc, ldw, t, piv_g, phase, pct0, pct1, nb1, nearly, next_rank, next_slot):
    cap = 1 << 16
    buf = (C.c_int32 * (5 * cap))()
    n = _native.host().jb_plan_enumerate(m, p, rank, rloc, ldw, t, piv_g % p, piv_g // p, int(piv_g != t), phase,
                                         pct0, pct1, nb1, nearly, next_rank, next_slot, buf, cap)
    assert n >= 0
    return [tuple(buf[5 * i + k] for k in range(5)) for i in range(n)]


class ScheduleModel:
    """Nr block rows of size m over p ranks (m = 128 unless given; other sizes make block columns
    straddle the 128-wide column tiles).  pivots[t] = global block row chosen at step t
    (>= t; default: the diagonal).  A pivot off the diagonal adds the block-row swap: ROWT pushes the
    old row t into the pivot owner's swap buffer, the swapped-in row is updated from there."""

    def __init__(self, Nr: int, p: int, sched: int = 2, b2_overlap: bool = False, k1_gated: bool = False,
                 frac: float = 0.3, count_rowt_in_early: bool = True, stepdone_handshake: bool = True,
                 pivots=None, swap_handshake: bool = True, mma_waits_early: bool = False, m: int = 128):
        self.Nr, self.p, self.m = Nr, p, m
        self.swap_handshake = swap_handshake
        self.mma_waits_early = mma_waits_early      # False = the shipped kernel (see _row_item)
        self.pivots = list(pivots) if pivots is not None else list(range(Nr))
        assert len(self.pivots) == Nr and all(t <= g < Nr for t, g in enumerate(self.pivots))
        self.sched, self.b2_overlap, self.k1_gated, self.frac = sched, b2_overlap, k1_gated, frac
        self.count_rowt_in_early, self.stepdone_handshake = count_rowt_in_early, stepdone_handshake
        from mpi_jordan_crazy_acceleration_b200.parallel.layout import padded_ld

        self.ldw = padded_ld(Nr * m, m)
        self.CT = (self.ldw + BN - 1) // BN    # block columns may straddle column tiles (m != 128)
        self.nsub = (m + 63) // 64
        self.g = Graph()
        self.flag: dict = {}               # flag key -> list of producer end nodes (published when all ended)
        self.waits: list = []              # (waiting node, flag key)
        self.acc = defaultdict(list)       # location -> [(start, end, is_write, label)]
        self.k1_start, self.k1_end = {}, {}
        for r in range(p):
            self._build_rank(r)
        self._resolve()

    # ------------------------------------------------------------------ helpers
    def rloc(self, r):
        return self.Nr // self.p + (1 if r < self.Nr % self.p else 0)

    def pri(self, t):
        return (t * self.m) // BN, min(self.CT - 1, ((t + 2) * self.m - 1) // BN)

    def split(self, t, last=False):
        pct0, pct1 = self.pri(t)
        nnp = self.CT - (pct1 - pct0 + 1)
        nearly = max(0, min(nnp, 8, self.pri(t + 1)[1] - pct1))
        nb1 = min(nnp, max(nearly, int(self.frac * nnp + 0.999)))
        if last:
            nb1 = nnp
        return pct0, pct1, nnp, nearly, nb1

    def tiles_of_block_col(self, t):
        return range((t * self.m) // BN, ((t + 1) * self.m - 1) // BN + 1)

    def np_pos(self, t, ct):
        pct0, pct1 = self.pri(t)
        return ct - pct1 - 1 if ct > pct1 else ct + self.CT - pct1 - 1

    def _access(self, loc, s, e, write, label):
        self.acc[loc].append((s, e, write, label))

    def _publish(self, key, node):
        self.flag.setdefault(key, []).append(node)

    # ------------------------------------------------------------------ kernels
    def _k1(self, stream: Stream, r: int, t: int, *preds):
        g = self.g
        s = stream.op(f"r{r} K1({t}) start", *preds)
        e = stream.op(f"r{r} K1({t}) end")
        self.k1_start[(r, t)], self.k1_end[(r, t)] = s, e
        lab = f"r{r} K1({t})"
        for slot in range(self.rloc(r)):
            for sub in range(self.nsub):
                for ct in self.tiles_of_block_col(t):                   # block column t of every local row
                    self._access(("W", r, slot, sub, ct), s, e, False, lab)
        self._access(("Lbuf", r, t % 3), s, e, True, lab)
        self._access(("invbuf", r, t & 1), s, e, True, lab)
        self._access(("stepinfo", r, t), s, e, True, lab)
        return e

    def _row_item(self, r, t, name, ks, ty, ct, lrow0, prow, wait_prev_early):
        """ROWT / UPD / UPD_SWAPPED item of step t; returns (end node, is ROWT)."""
        g, p, m = self.g, self.p, self.m
        piv_g = self.pivots[t]
        piv_rank, piv_slot, do_swap = piv_g % p, piv_g // p, piv_g != t
        slot, sub = lrow0 // m, (lrow0 % m) // 64
        kind = {ROWT: "ROWT", UPD: "UPDP" if prow else "UPD", UPD_SW: "UPDSW"}[ty]
        lab = f"r{r} {name}({t}) {kind} slot{slot} sub{sub} ct{ct}"
        s = g.node(lab + " start", ks)
        if wait_prev_early is not None:
            self.waits.append((s, wait_prev_early))
        tiles_t = set(self.tiles_of_block_col(t))   # column tiles holding block column t
        if ty == ROWT:
            cur = s
            if do_swap:                             # push the old row t to the pivot owner (main.cpp:1100-1131)
                if t >= 2 and self.stepdone_handshake and self.swap_handshake:
                    self.waits.append((s, ("stepdone", piv_rank, t - 2)))
                mid = g.node(lab + " pushed", s)
                self._access(("W", r, slot, sub, ct), s, mid, False, lab + " push")
                self._access(("swapbuf", piv_rank, t & 1, ct, sub), s, mid, True, lab + " push")
                self._publish(("swapflag", t, ct, sub), mid)
                cur = mid
            s2 = g.node(lab + " install", cur)
            self.waits.append((s2, ("rowflag", t, ct)))
            e = g.node(lab + " end", s2)
            self._access(("stepinfo", r, t), s, e, False, lab)
            self._access(("rowbuf", r, t & 1, ct, sub), s2, e, False, lab)
            self._access(("W", r, slot, sub, ct), s2, e, True, lab)
            return e, True
        e = g.node(lab + " end", s)
        self.waits.append((s, ("rowflag", t, ct)))
        # The flag waits above are made by the TMA producer thread and order the operand loads.  The
        # MMA warps read the C tile straight from W as soon as the item has been handed to them, i.e.
        # possibly BEFORE those waits are over (csrc/step_fused.cu: the ring slot is published before
        # the producer waits).  Only the swapped-in row's MMA warps wait themselves (swap flag).
        c_start = s if (ty == UPD_SW or self.mma_waits_early) else g.node(lab + " C-read", ks)
        if c_start is not s:
            g.edge(c_start, e)
        self._access(("stepinfo", r, t), s, e, False, lab)
        for sb in range(self.nsub):
            self._access(("rowbuf", r, t & 1, ct, sb), s, e, False, lab)
        if ty == UPD_SW:
            assert do_swap and r == piv_rank and slot == piv_slot
            for c in tiles_t | {ct}:
                self.waits.append((s, ("swapflag", t, c, sub)))
                self._access(("swapbuf", r, t & 1, c, sub), s, e, False, lab)
        else:
            self._access(("Lbuf", r, t % 3), s, e, False, lab)
            self._access(("W", r, slot, sub, ct), c_start, e, False, lab + " C-read")
        self._access(("W", r, slot, sub, ct), s, e, True, lab)
        if prow:
            self._publish(("prow", r, t, ct, sub), e)
        return e, False

    def _step_kernel(self, stream: Stream, r: int, t: int, phase: int, last=False, *preds):
        """One launch of k_step_fused; returns its end node."""
        g, p, m = self.g, self.p, self.m
        pct0, pct1, nnp, nearly, nb1 = self.split(t, last)
        nxt = self.pivots[t + 1] if (phase == PH_B2 and t + 1 < self.Nr) else -1
        items = _enumerate(m, p, r, self.rloc(r), self.ldw, t, self.pivots[t], phase, pct0, pct1, nb1, nearly,
                           nxt % p if nxt >= 0 else -1, nxt // p if nxt >= 0 else 0)
        name = {PH_A: "A", PH_S: "S", PH_B1: "B1", PH_B2: "B2"}[phase]
        ks = stream.op(f"r{r} {name}({t}) start", *preds)
        # every CTA reads stepinfo[t] (B2: also stepinfo[t+1]) when it starts, before any flag wait
        self._access(("stepinfo", r, t), ks, ks, False, f"r{r} {name}({t}) prologue")
        if nxt >= 0:
            self._access(("stepinfo", r, t + 1), ks, ks, False, f"r{r} {name}({t}) prologue")
        ends = []
        c0 = nb1 if phase == PH_B2 else 0
        early_items = defaultdict(list)
        for (ty, ct, lrow0, rows, prow) in items:
            if ty == SCALE:
                st = t + 1 if phase == PH_B2 else t          # step whose pivot row is produced
                sub = lrow0 // 64
                lab = f"r{r} {name}({t}) SCALE({st}) ct{ct} sub{sub}"
                s = g.node(lab + " start", ks)
                e = g.node(lab + " end", s)
                if phase == PH_B2:
                    if t > 0 and self.stepdone_handshake:
                        for q in range(p):
                            self.waits.append((s, ("stepdone", q, t - 1)))
                    if not (pct0 <= ct <= pct1):
                        pos = self.np_pos(t, ct)
                        if pos >= c0:
                            for sb in range(self.nsub):
                                self.waits.append((s, ("prow", r, t, ct, sb)))
                        elif pos < nearly:
                            self.waits.append((s, ("early", r, t, pos)))
                        else:
                            self.waits.append((s, ("b1done", r, t)))
                assert self.pivots[st] % p == r
                pslot = self.pivots[st] // p
                for sb in range(self.nsub):
                    self._access(("W", r, pslot, sb, ct), s, e, False, lab)
                self._access(("invbuf", r, st & 1), s, e, False, lab)
                self._access(("stepinfo", r, st), s, e, False, lab)
                for q in range(p):
                    self._access(("rowbuf", q, st & 1, ct, sub), s, e, True, lab)
                self._publish(("rowflag", st, ct), e)
            else:
                prev_early = None
                if phase == PH_A and t > 0:
                    ppct0, ppct1 = self.pri(t - 1)
                    if ct < ppct0 or ct > ppct1:
                        j = ct - ppct1 - 1
                        if j < 0:
                            j += self.CT
                        if j < 8:
                            prev_early = ("early", r, t - 1, j)
                e, is_rowt = self._row_item(r, t, name, ks, ty, ct, lrow0, prow, prev_early)
                if phase == PH_B1 and not (pct0 <= ct <= pct1):
                    pos = self.np_pos(t, ct)
                    if pos < nearly and (not is_rowt or self.count_rowt_in_early):
                        early_items[pos].append(e)
            ends.append(e)
        ke = stream.op(f"r{r} {name}({t}) end", *ends)
        if phase == PH_B1:
            for pos in range(nearly):
                fl = g.node(f"r{r} early({t},{pos})", ks, *early_items.get(pos, []))
                self._publish(("early", r, t, pos), fl)
            self._publish(("b1done", r, t), ke)
        if phase == PH_B2:
            sd = g.node(f"r{r} stepdone({t})", ke)
            self.waits.append((sd, ("b1done", r, t)))
            self._publish(("stepdone", r, t), sd)
        return ke

    # ------------------------------------------------------------------ host sequence (jb_engine_invert)
    def _build_rank(self, r: int):
        g, Nr = self.g, self.Nr
        st, hs, bs = Stream(g, f"r{r}.st"), Stream(g, f"r{r}.h"), Stream(g, f"r{r}.b")
        ev = {}

        def record(name, stream):
            ev[name] = stream.op(f"record {name}")

        def wait(stream, name):
            stream.op(f"wait {name}", ev[name])

        self._k1(st, r, 0)
        self._step_kernel(st, r, 0, PH_S)
        if self.sched == 1:          # phase A in line: hstream A(t) -> K1(t+1); st B1(t); bstream B2(t)
            record("B1[1]", st)
            record("B2[1]", st)
            for t in range(Nr - 1):
                wait(hs, f"B1[{(t + 1) & 1}]")
                wait(hs, f"B2[{(t + 1) & 1}]")
                self._step_kernel(hs, r, t, PH_A)
                record(f"G[{t & 1}]", hs)
                self._k1(hs, r, t + 1)
                record(f"K[{(t + 1) & 1}]", hs)
                wait(st, f"G[{t & 1}]")
                self._step_kernel(st, r, t, PH_B1)
                record(f"B1[{t & 1}]", st)
                wait(bs, f"K[{(t + 1) & 1}]")
                if not self.b2_overlap:
                    wait(bs, f"B1[{t & 1}]")
                self._step_kernel(bs, r, t, PH_B2)
                record(f"B2[{t & 1}]", bs)
            t = Nr - 1
            wait(st, f"B2[{(t + 1) & 1}]")
            self._wide_last(st, r, t)
            return
        # sched 2: phase A(t+1) beside B2(t), K1(t+2) pending behind it
        record("G[1]", st)
        wait(hs, "G[1]")
        self._step_kernel(hs, r, 0, PH_A)
        record("A", hs)
        self._k1(hs, r, 1)
        record("K[1]", hs)
        wait(st, "A")
        wait(bs, "G[1]")
        for t in range(Nr - 1):
            if t > 0:
                wait(st, f"B2[{(t + 1) & 1}]")
            self._step_kernel(st, r, t, PH_B1)
            record(f"B1[{t & 1}]", st)
            wait(bs, f"K[{(t + 1) & 1}]")
            if t > 0:
                wait(bs, f"B1[{(t + 1) & 1}]")
            if not self.b2_overlap:
                wait(bs, f"B1[{t & 1}]")
            record(f"G[{t & 1}]", bs)
            self._step_kernel(bs, r, t, PH_B2)
            record(f"B2[{t & 1}]", bs)
            wait(hs, f"G[{t & 1}]")
            self._step_kernel(hs, r, t + 1, PH_A)
            if t + 2 < Nr:
                if self.k1_gated:
                    wait(hs, f"B2[{t & 1}]")
                self._k1(hs, r, t + 2)
                record(f"K[{t & 1}]", hs)
            else:
                record("A", hs)
        t = Nr - 1
        wait(st, f"B2[{(t + 1) & 1}]")
        wait(st, "A")
        self._step_kernel(st, r, t, PH_B1, True)

    def _wide_last(self, stream, r, t):
        """Last step of the in-line schedule: phase A over every column tile (ROWT + UPD of everything)."""
        items = _enumerate(self.m, self.p, r, self.rloc(r), self.ldw, t, self.pivots[t], PH_A, 0, self.CT - 1, 0, 0,
                           -1, 0)
        ks = stream.op(f"r{r} Awide({t}) start")
        ends = [self._row_item(r, t, "Awide", ks, ty, ct, lrow0, prow, None)[0] for (ty, ct, lrow0, rows, prow) in items]
        stream.op(f"r{r} Awide({t}) end", *ends)

    # ------------------------------------------------------------------ cross-rank edges + check
    def _resolve(self):
        g = self.g
        for node, key in self.waits:
            prods = self.flag.get(key)
            assert prods, f"nobody publishes {key} (waited for by {g.names[node]})"
            for pnode in prods:
                g.edge(pnode, node)
        for (r, t), e in self.k1_end.items():      # min-loc: every rank's candidate must have been published
            for q in range(self.p):
                g.edge(self.k1_start[(q, t)], e)
        self.reach = g.closure()

    def before(self, a_end: int, b_start: int) -> bool:
        return a_end == b_start or bool((self.reach[b_start] >> a_end) & 1)

    def hazards(self, limit: int = 20) -> list[str]:
        out = []
        for loc, lst in self.acc.items():
            writes = [x for x in lst if x[2]]
            for w in writes:
                for o in lst:
                    if o is w or o[1] == w[1] or (o[2] and id(o) < id(w)):   # same item / pair seen from the other side
                        continue
                    if self.before(w[1], o[0]) or self.before(o[1], w[0]):
                        continue
                    out.append(f"{loc}: [{w[3]}] (write) unordered with [{o[3]}] ({'write' if o[2] else 'read'})")
                    if len(out) >= limit:
                        return out
        return out
