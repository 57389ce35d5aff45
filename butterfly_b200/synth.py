"""Synthetic flattened operators with the SHAPE of what the reference's factorizers produce
(src/fac_helm2.c:653-704, 806-1002): a HODLR-style hierarchy whose near-field blocks are dense
leaves and whose well-separated blocks are complementary butterflies — MatProduct [BlockDiag
(last), BlockCoo x (L-2), BlockDiag (first)] with ragged ranks.  Used by smoke(), by parity tests
(random values, small n) and by bench.py (value-less plans at the named size; the pool is then
filled on the device by bfCudaPlanFillSynthetic).

This is input generation, not part of the apply path.
"""
from __future__ import annotations

import numpy as np

from . import capi
from .rawops import RawOps


class _Builder:
    def __init__(self, dtype, with_values, rng):
        self.dtype = dtype
        self.cplx = dtype == capi.COMPLEX_F64
        self.with_values = with_values
        self.rng = rng
        self.ops = []       # tuples (kind, flags, inBuf, inOff, outBuf, outOff, m, k, ld, valOff)
        self.vals = []      # float64 chunks
        self.nvals = 0
        self.buf_rows = [0, 0]

    def new_buf(self, rows):
        self.buf_rows.append(int(rows))
        return len(self.buf_rows) - 1

    def _push_vals(self, a):
        off = self.nvals
        a = np.ascontiguousarray(a, dtype=np.float64).reshape(-1)
        self.vals.append(a)
        self.nvals += a.size
        return off

    def dense(self, in_buf, in_off, out_buf, out_off, m, k, trans=False, pad=0):
        """acting shape m x k; trans: stored as the (conjugate-)transposed k x m matrix with the
        TRANS (and, for complex, CONJ) property bits, as bfMatTranspose leaves it
        (src/mat_dense_complex.c:1475-1478); pad: extra row stride (a view into a wider parent)."""
        flags = 0
        sr, sc = (k, m) if trans else (m, k)
        if trans:
            flags = capi.LEAF_TRANS | (capi.LEAF_CONJ if self.cplx else 0)
        ld = sc + pad
        off = 0
        if self.with_values:
            scale = 1.0 / np.sqrt(max(k, 1))
            e = 2 if self.cplx else 1
            a = self.rng.standard_normal((sr, ld * e)) * scale
            off = self._push_vals(a)
        self.ops.append((capi.LEAF_DENSE, flags, in_buf, in_off, out_buf, out_off, m, k, ld, off))

    def identity(self, in_buf, in_off, out_buf, out_off, m, k):
        self.ops.append((capi.LEAF_IDENTITY, 0, in_buf, in_off, out_buf, out_off, m, k, 0, 0))

    def diag(self, in_buf, in_off, out_buf, out_off, m, k):
        off = 0
        if self.with_values:
            off = self._push_vals(0.5 + self.rng.random(min(m, k)))
        self.ops.append((capi.LEAF_DIAG, 0, in_buf, in_off, out_buf, out_off, m, k, 1, off))

    def finish(self, num_rows, num_cols) -> RawOps:
        self.buf_rows[0] = num_cols
        self.buf_rows[1] = num_rows
        ops = np.zeros(len(self.ops), capi.RAWOP_DTYPE)
        for i, name in enumerate(("kind", "flags", "inBuf", "inOff", "outBuf", "outOff", "m", "k", "ld", "val")):
            ops[name] = [o[i] for o in self.ops]
        vals = np.concatenate(self.vals) if (self.with_values and self.vals) else (np.zeros(0) if self.with_values else None)
        return RawOps(self.dtype, num_rows, num_cols, np.array(self.buf_rows, np.int64), ops, vals)


def _split(lo, hi, depth):
    """Leaves of a balanced binary tree of the given depth over [lo, hi)."""
    edges = np.linspace(lo, hi, (1 << depth) + 1).round().astype(np.int64)
    return [(int(edges[i]), int(edges[i + 1])) for i in range(1 << depth)]


def _butterfly(b: _Builder, in_buf, src, out_buf, tgt, depth, rank_lo, rank_hi, exotic):
    """Complementary butterfly of `depth` inner levels between source rows src = (lo, hi) of in_buf
    and target rows tgt of out_buf.  Level l pairs target nodes of depth l with source nodes of
    depth (depth - l); the pair (t, s) carries rank r[l][t][s]."""
    rng = b.rng
    d = depth
    r = [rng.integers(rank_lo, rank_hi + 1, size=(1 << l, 1 << (d - l))) for l in range(d + 1)]
    src_leaves = _split(src[0], src[1], d)
    tgt_leaves = _split(tgt[0], tgt[1], d)
    # factor applied first: BlockDiag over source leaves (src/fac_helm2.c:42-220)
    offs = np.concatenate([[0], np.cumsum(r[0][0, :])])
    cur = b.new_buf(int(offs[-1]))
    for s, (lo, hi) in enumerate(src_leaves):
        b.dense(in_buf, lo, cur, int(offs[s]), int(r[0][0, s]), hi - lo, trans=exotic and s % 3 == 1, pad=(2 if exotic and s % 4 == 2 else 0))
    cur_off = {(0, s): int(offs[s]) for s in range(1 << d)}
    # inner factors: BlockCoo, one block per (target child, source child) (src/fac_helm2.c:222-401)
    for l in range(1, d + 1):
        nt, ns = 1 << l, 1 << (d - l)
        sizes = r[l].reshape(-1)
        noffs = np.concatenate([[0], np.cumsum(sizes)])
        nxt = b.new_buf(int(noffs[-1]))
        nxt_off = {}
        for t in range(nt):
            for s in range(ns):
                o = int(noffs[t * ns + s])
                nxt_off[(t, s)] = o
        # storage order of a BlockCoo is not row-sorted in the reference: shuffle it
        blocks = [(t, s, c) for t in range(nt) for s in range(ns) for c in (0, 1)]
        if exotic:
            rng.shuffle(blocks)
        for t, s, c in blocks:
            tp, sc = t >> 1, 2 * s + c
            b.dense(cur, cur_off[(tp, sc)], nxt, nxt_off[(t, s)], int(r[l][t, s]), int(r[l - 1][tp, sc]))
        cur, cur_off = nxt, nxt_off
    # factor applied last: BlockDiag over target leaves (src/fac_helm2.c:403-549)
    for t, (lo, hi) in enumerate(tgt_leaves):
        b.dense(cur, cur_off[(t, 0)], out_buf, lo, hi - lo, int(r[d][t, 0]))


def _hodlr(b: _Builder, lo, hi, leaf, rank_lo, rank_hi, exotic, level=0):
    n = hi - lo
    if n <= 2 * leaf:
        b.dense(0, lo, 1, lo, n, n)  # near-field dense leaf (src/fac_helm2.c:888)
        return
    mid = (lo + hi) // 2
    halves = [(lo, mid), (mid, hi)]
    for i in (0, 1):
        for j in (0, 1):
            if i == j:
                _hodlr(b, halves[i][0], halves[i][1], leaf, rank_lo, rank_hi, exotic, level + 1)
            else:
                size = min(halves[i][1] - halves[i][0], halves[j][1] - halves[j][0])
                depth = max(0, int(np.log2(max(size // leaf, 1))))
                _butterfly(b, 0, halves[j], 1, halves[i], depth, rank_lo, rank_hi, exotic)


def random_hodbf(n=2048, dtype=capi.COMPLEX_F64, seed=0, leaf=48, rank_lo=9, rank_hi=23, with_values=True,
                 exotic=True) -> RawOps:
    """A square n x n HODLR-of-butterflies operator.  exotic=True adds what the reference can
    contain besides plain dense leaves: transposed views, padded row strides, shuffled BlockCoo
    storage order, an identity and a real-diagonal term (MatSum with 0.5 I,
    examples/multiple_scattering/multiple_scattering_context.c:483-487) and an uncovered row gap."""
    rng = np.random.default_rng(seed)
    b = _Builder(dtype, with_values, rng)
    _hodlr(b, 0, n, leaf, rank_lo, rank_hi, exotic)
    if exotic:
        b.diag(0, 0, 1, 0, n, n)
        q = n // 4
        b.identity(0, q, 1, 2 * q, q, q)
    return b.finish(n, n)


def random_rect(num_rows=700, num_cols=300, dtype=capi.REAL_F64, seed=0, with_values=True) -> RawOps:
    """A rectangular MatProduct [BlockDiag, BlockDiag] with an uncovered output gap and an
    identity pass-through block, the shape of a streamed (LBO) partial factorization
    (src/fac.c:649-676)."""
    rng = np.random.default_rng(seed)
    b = _Builder(dtype, with_values, rng)
    ncb = 5
    cedges = np.linspace(0, num_cols, ncb + 1).round().astype(int)
    ranks = rng.integers(7, 30, ncb)
    offs = np.concatenate([[0], np.cumsum(ranks)])
    t = b.new_buf(int(offs[-1]))
    for j in range(ncb):
        k = int(cedges[j + 1] - cedges[j])
        if j == 2:
            ranks_j = min(int(ranks[j]), k)
            b.identity(0, int(cedges[j]), t, int(offs[j]), ranks_j, ranks_j)  # pass-through block
        else:
            b.dense(0, int(cedges[j]), t, int(offs[j]), int(ranks[j]), k, trans=(j == 1))
    redges = np.linspace(0, num_rows - 37, ncb + 1).round().astype(int)  # last 37 rows: nothing writes them
    for j in range(ncb):
        b.dense(t, int(offs[j]), 1, int(redges[j]), int(redges[j + 1] - redges[j]), int(ranks[j]))
    return b.finish(num_rows, num_cols)


def lbo_like(num_rows=500_000, num_eigs=4096, bands=16, depth=5, rank_lo=20, rank_hi=44, dtype=capi.REAL_F64,
             seed=0, with_values=False) -> RawOps:
    """Shape of a streamed LBO eigenbasis butterfly (bfFacSpanGetMat, src/fac_span.c:126-155): a
    1 x bands MatBlockDense whose blocks are MatProducts mapping one frequency band (a contiguous
    run of eigenvector columns) to ALL vertices through `depth` butterfly levels.  Structure only
    (values filled on the device) unless with_values; stands in for configs[3] at sizes the
    reference's CPU streamer cannot reach in bench time (169 s already at 4000 x 1024)."""
    rng = np.random.default_rng(seed)
    b = _Builder(dtype, with_values, rng)
    edges = np.linspace(0, num_eigs, bands + 1).round().astype(np.int64)
    for j in range(bands):
        _butterfly(b, 0, (int(edges[j]), int(edges[j + 1])), 1, (0, num_rows), depth, rank_lo, rank_hi, False)
    return b.finish(num_rows, num_eigs)


def shifted_identity_hodbf(n=1024, seed=5, scale=0.005) -> RawOps:
    """I + (random exotic HODBF with its dense values scaled by `scale`): a well-conditioned complex
    system for the GMRES tests (converges in a few dozen steps)."""
    raw = random_hodbf(n=n, dtype=capi.COMPLEX_F64, seed=seed, exotic=True)
    raw.values = raw.values * scale
    ops = np.concatenate([raw.ops, np.zeros(1, raw.ops.dtype)])
    o = ops[-1:]
    o["kind"], o["flags"], o["inBuf"], o["inOff"], o["outBuf"], o["outOff"] = capi.LEAF_IDENTITY, 0, 0, 0, 1, 0
    o["m"], o["k"], o["ld"], o["val"] = raw.num_rows, raw.num_cols, 0, 0
    return RawOps(raw.dtype, raw.num_rows, raw.num_cols, raw.buf_rows, ops, raw.values)
