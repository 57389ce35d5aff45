"""Block-row sharding of one operator across ranks (SURVEY.md §8e, configs[4]).

Every rank holds the SAME structure (all leaves: identical stages, workspace layout and row
addresses everywhere) but applies — and stores the factor values of — only the leaves whose output
rows it owns; the others carry BF_CUDA_LEAF_REMOTE.  Ownership is by contiguous row ranges:

  * Y is cut into `world` ranges of (nearly) equal rows, snapped to rows no leaf straddles;
  * a MatProduct chain (its intermediate buffers) whose Y rows lie inside one range lives entirely on
    that rank — no exchange at all (every near-field leaf and most butterflies of the HODBF);
  * a chain whose Y rows span several ranks has each intermediate buffer cut in the same proportions.
    Early buffers are ordered by source box, late ones by target box (src/fac_helm2.c:222-401), so a
    rank's rows of level l mostly depend on its own rows of level l - 1; what does not is exchanged.

After stage t each rank sends the workspace rows it produced at t that another rank reads later, in
ONE all_to_all_single (NCCL over NVLink on GPUs, gloo in the CPU tests).  For a butterfly this is the
mid-level exchange of the north star: the early levels are source-local, the late ones target-local,
and the rows that change hands concentrate where the ordering flips.  X is replicated (it is n x nrhs,
tiny next to the factors); every rank ends up with its own rows of Y.

Host logic + torch.distributed plumbing only; the stage kernels run through bfCudaApplyStages.
"""
from __future__ import annotations

import bisect
import ctypes as C
from typing import List

import numpy as np

from . import capi
from .rawops import RawOps

LEAF_REMOTE = 8  # BF_CUDA_LEAF_REMOTE


# ---- ownership ------------------------------------------------------------------------------------
def op_stages(raw: RawOps) -> np.ndarray:
    """Stage of every raw op in the forward direction (the rule of compileDirection, bf_plan.cpp §2)."""
    level = np.zeros(len(raw.buf_rows), np.int64)
    st = np.zeros(raw.num_ops, np.int64)
    ib, ob = raw.ops["inBuf"], raw.ops["outBuf"]
    for i in range(raw.num_ops):
        s = level[ib[i]]
        st[i] = s
        if level[ob[i]] < s + 1:
            level[ob[i]] = s + 1
    return st


def _clean_cuts(starts, ends):
    """Rows where a buffer can be cut without splitting a leaf: sorted unique starts s with every
    earlier leaf ending at or before s."""
    order = np.argsort(starts, kind="stable")
    s, e = starts[order], ends[order]
    run = np.maximum.accumulate(np.concatenate([[0], e[:-1]]))
    ok = s >= run
    return np.unique(s[ok])


def _cut_rows(starts, ends, rows_total, fractions):
    """Cut [0, rows_total) at the clean cuts closest to the cumulative `fractions` (len = parts).
    Returns parts + 1 nondecreasing boundaries."""
    cand = np.concatenate([_clean_cuts(starts, ends), [rows_total]]) if len(starts) else np.array([rows_total])
    cum = np.cumsum(fractions) / np.sum(fractions)
    b = [0]
    for f in cum[:-1]:
        target = f * rows_total
        j = int(np.argmin(np.abs(cand - target)))
        b.append(max(int(cand[j]), b[-1]))
    b.append(int(rows_total))
    return b


def assign_owners(raw: RawOps, y_cuts) -> np.ndarray:
    """owner[i] = rank that applies raw op i, given the Y row boundaries (no leaf writing Y may
    straddle one: shard() arranges that)."""
    from scipy.sparse import coo_matrix
    from scipy.sparse.csgraph import connected_components

    ops = raw.ops
    n_ops = raw.num_ops
    ib, ob = ops["inBuf"].astype(np.int64), ops["outBuf"].astype(np.int64)
    oo, m = ops["outOff"].astype(np.int64), ops["m"].astype(np.int64)
    cuts = np.asarray(y_cuts, np.int64)
    nparts = len(cuts) - 1
    row_owner = lambda r: np.minimum(np.searchsorted(cuts, r, side="right") - 1, nparts - 1)
    owner = np.full(n_ops, -1, np.int64)

    # chains: connected components of intermediate buffers (a MatProduct's temporaries)
    nb = len(raw.buf_rows)
    both = (ib >= 2) & (ob >= 2)
    g = coo_matrix((np.ones(int(both.sum()), np.int8), (ib[both], ob[both])), shape=(nb, nb))
    _, comp = connected_components(g, directed=False)
    temp = np.where(ob >= 2, ob, np.where(ib >= 2, ib, -1))
    chain = np.where(temp >= 0, comp[np.maximum(temp, 0)], -1)
    direct = chain < 0
    owner[direct] = row_owner(oo[direct])  # leaves that read X and write Y directly

    idx = np.nonzero(~direct)[0]
    if len(idx) == 0:
        return owner
    order = idx[np.argsort(chain[idx], kind="stable")]
    cid = chain[order]
    starts = np.nonzero(np.concatenate([[True], cid[1:] != cid[:-1]]))[0]
    ends = np.concatenate([starts[1:], [len(order)]])
    # Y extent of every chain (leaves with outBuf == 1), vectorised
    is_last = ob[order] == 1
    big = np.iinfo(np.int64).max
    lo = np.minimum.reduceat(np.where(is_last, oo[order], big), starts)
    hi = np.maximum.reduceat(np.where(is_last, oo[order] + m[order], -1), starts)
    never = hi < 0  # a chain that never reaches Y (degenerate): rank 0
    p0 = np.where(never, 0, row_owner(np.where(never, 0, lo)))
    p1 = np.where(never, 0, row_owner(np.maximum(np.where(never, 0, hi) - 1, np.where(never, 0, lo))))
    chain_len = ends - starts
    single = p0 == p1
    owner[order] = np.repeat(np.where(single, p0, -1), chain_len)
    for c in np.nonzero(~single)[0]:  # chains whose Y rows span ranks: cut every buffer proportionally
        sel = order[starts[c]:ends[c]]
        r0, r1 = int(lo[c]), int(hi[c])
        ranks = list(range(int(p0[c]), int(p1[c]) + 1))
        frac = np.array([max(min(int(cuts[p + 1]), r1) - max(int(cuts[p]), r0), 0) for p in ranks], np.float64)
        last = sel[ob[sel] == 1]
        owner[last] = row_owner(oo[last])
        tsel = sel[ob[sel] >= 2]
        for bb in np.unique(ob[tsel]):
            w = tsel[ob[tsel] == bb]
            bc = np.asarray(_cut_rows(oo[w], oo[w] + m[w], int(raw.buf_rows[bb]), frac), np.int64)
            owner[w] = np.asarray(ranks)[np.minimum(np.searchsorted(bc, oo[w], side="right") - 1, len(ranks) - 1)]
    assert (owner >= 0).all()
    return owner


def _split_op_rows(o, bounds, dtype, has_values):
    """Slabs of one raw op cut at the acting-row offsets in `bounds` (relative, ascending, inside
    (0, m)).  Row-major storage: a slab of acting rows [r, r + h) starts r * ld elements into the
    stored matrix, or r elements for a TRANS-stored leaf (acting rows are stored columns); a diagonal
    slab starts r doubles in.  Identity / diagonal leaves act on min(m, k) rows."""
    e = 2 if dtype == capi.COMPLEX_F64 else 1
    m = int(o["m"])
    kind = int(o["kind"])
    stored_t = bool(int(o["flags"]) & (capi.LEAF_TRANS | (capi.LEAF_CONJ if dtype == capi.COMPLEX_F64 else 0)))
    edges = [0] + [int(b) for b in bounds if 0 < b < m] + [m]
    out = []
    for r, r1 in zip(edges, edges[1:]):
        h = r1 - r
        c = o.copy()
        c["outOff"] = int(o["outOff"]) + r
        c["m"] = h
        if kind != capi.LEAF_DENSE:
            if r >= int(o["k"]):
                break
            c["inOff"] = int(o["inOff"]) + r
            c["k"] = min(h, int(o["k"]) - r)
            c["m"] = c["k"]
            if kind == capi.LEAF_DIAG and has_values:
                c["val"] = int(o["val"]) + r
        elif has_values:
            c["val"] = int(o["val"]) + (r * e if stored_t else r * int(o["ld"]) * e)
        out.append(c)
    return out


def origin_of(raw: RawOps):
    """(op index, acting-row offset) of every op of `raw` in the op list shard() was given (identity when
    nothing was cut): what maps per-leaf side data — the Helmholtz recipes of helm2.Recording — through
    the cuts."""
    o = getattr(raw, "origin", None)
    return o if o is not None else (np.arange(raw.num_ops, dtype=np.int64), np.zeros(raw.num_ops, np.int64))


def _with_origin(new: RawOps, old: RawOps, src) -> RawOps:
    src = np.asarray(src, dtype=np.int64)
    oi, oo = origin_of(old)
    off = new.ops["outOff"].astype(np.int64) - old.ops["outOff"].astype(np.int64)[src]
    new.origin = (oi[src], oo[src] + off)
    return new


def presplit_tall_leaves(raw: RawOps, max_rows: int = 2048) -> RawOps:
    """Cut dense leaves taller than max_rows into slabs so ownership boundaries have somewhere to fall
    (e.g. the 15k-row last factors of a streamed eigenbasis).  Summation order per row is unchanged."""
    ops = raw.ops
    tall = (ops["m"] > max_rows) & (ops["kind"] == capi.LEAF_DENSE)
    if not tall.any():
        return raw
    out, src = [], []
    for i in range(raw.num_ops):
        if tall[i]:
            parts = _split_op_rows(ops[i], range(max_rows, int(ops[i]["m"]), max_rows), raw.dtype, raw.values is not None)
        else:
            parts = [ops[i].copy()]
        out.extend(parts)
        src.extend([i] * len(parts))
    return _with_origin(RawOps(raw.dtype, raw.num_rows, raw.num_cols, raw.buf_rows, np.array(out, dtype=ops.dtype), raw.values), raw, src)


def split_diagonal_terms(raw: RawOps, y_cuts) -> RawOps:
    """Identity / diagonal leaves that write Y (the n x n "I / 2" MatSum term of
    multiple_scattering_context.c:483-487) can be cut anywhere: cut them at the Y boundaries."""
    ops = raw.ops
    sel = (ops["kind"] != capi.LEAF_DENSE) & (ops["outBuf"] == 1)
    if not sel.any():
        return raw
    out, src = [], []
    for i in range(raw.num_ops):
        if sel[i]:
            o0 = int(ops[i]["outOff"])
            parts = _split_op_rows(ops[i], [c - o0 for c in y_cuts], raw.dtype, raw.values is not None)
        else:
            parts = [ops[i].copy()]
        out.extend(parts)
        src.extend([i] * len(parts))
    return _with_origin(RawOps(raw.dtype, raw.num_rows, raw.num_cols, raw.buf_rows, np.array(out, dtype=ops.dtype), raw.values), raw, src)


def choose_y_cuts(raw: RawOps, world: int, fractions=None) -> List[int]:
    """world + 1 row boundaries of Y: equal shares (or `fractions`) snapped to rows no DENSE leaf straddles."""
    ops = raw.ops
    wy = np.nonzero((ops["outBuf"] == 1) & (ops["kind"] == capi.LEAF_DENSE))[0]
    oo, m = ops["outOff"].astype(np.int64), ops["m"].astype(np.int64)
    return _cut_rows(oo[wy], oo[wy] + m[wy], raw.num_rows, np.ones(world) if fractions is None else np.asarray(fractions, np.float64))


def shard(raw: RawOps, world: int, fractions=None):
    """(raw', owner, y_cuts): the operator with tall / diagonal leaves cut at ownership boundaries, the
    rank that applies each of its leaves, and the Y row ranges."""
    raw = presplit_tall_leaves(raw)
    y_cuts = choose_y_cuts(raw, world, fractions)
    raw = split_diagonal_terms(raw, y_cuts)
    owner = assign_owners(raw, y_cuts)
    return raw, owner, y_cuts


def mask_for_rank(raw: RawOps, owner: np.ndarray, rank: int) -> RawOps:
    """The rank's view of the operator: same leaves, those of other ranks flagged REMOTE."""
    ops = raw.ops.copy()
    remote = owner != rank
    ops["flags"][remote] |= LEAF_REMOTE
    if raw.values is not None:
        ops["val"][remote & (ops["kind"] == capi.LEAF_IDENTITY)] = 0
    out = RawOps(raw.dtype, raw.num_rows, raw.num_cols, raw.buf_rows, ops, raw.values)
    out.origin = getattr(raw, "origin", None)
    out.keep_leaves = True  # every rank must compile the SAME stages and buffers, also the one that owns every leaf
    return out


# ---- exchange lists -------------------------------------------------------------------------------
def _runs(mask: np.ndarray) -> np.ndarray:
    """(k, 2) array of [start, end) runs of True."""
    d = np.diff(np.concatenate([[0], mask.view(np.int8), [0]]))
    return np.stack([np.nonzero(d == 1)[0], np.nonzero(d == -1)[0]], axis=1)


def _mask_from_runs(runs, n):
    d = np.zeros(n + 1, np.int8)
    if len(runs):
        d[runs[:, 0]] += 1
        d[runs[:, 1]] -= 1  # runs are disjoint and sorted: the prefix sum is 0 / 1
    return np.cumsum(d[:-1], dtype=np.int8).astype(bool)


class StageTables:
    """What the exchange needs from one rank's compiled tables: per stage, the workspace rows its
    items write (have) and its tiles read (need)."""

    def __init__(self, raw_masked: RawOps, num_queues: int):
        lib = capi.load_library()
        desc = raw_masked.desc()
        t = C.POINTER(capi.DebugTables)()
        capi.check(lib.bfCudaDebugCompile(C.byref(desc), capi.OP_N, num_queues, C.byref(t)), "bfCudaDebugCompile")
        try:
            d = t.contents
            nS = d.numStages
            arr = lambda p, n, dt: np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_ubyte)), shape=(n * np.dtype(dt).itemsize,)).view(dt).copy() if n else np.zeros(0, dt)
            si = arr(d.stageItemBegin, nS + 1, np.int64)
            sti = arr(d.stageTileBegin, nS + 1, np.int64)
            items = arr(d.items, d.numItems, capi.ITEM_DTYPE)
            tiles = arr(d.tiles, d.numTiles, capi.TILE_DTYPE)
            self.num_stages = nS
            self.ws_rows = int(d.wsRows)
            self.arena = (int(d.wsArena0Rows), int(d.wsArena1Rows))
            self.have, self.need = [], []
            for s in range(nS):
                it = items[si[s]:si[s + 1]]
                sp = it["outAddr"] >> 30
                w = it[sp == capi.SPACE_W]
                hm = np.zeros(self.ws_rows + 1, np.int32)
                r0 = (w["outAddr"] & capi.ROW_MASK).astype(np.int64)
                np.add.at(hm, r0, 1)
                np.add.at(hm, r0 + w["rows"].astype(np.int64), -1)
                self.have.append(_runs(np.cumsum(hm)[:-1] > 0))
                tl = tiles[sti[s]:sti[s + 1]]
                tw = tl[(tl["inAddr"] >> 30) == capi.SPACE_W]
                kc = tw["kc"].astype(np.int64)
                k_eff = np.where(kc >= capi.KC_DIAG, tw["rows"].astype(np.int64), kc)
                nm = np.zeros(self.ws_rows + 1, np.int32)
                t0 = (tw["inAddr"] & capi.ROW_MASK).astype(np.int64)
                np.add.at(nm, t0, 1)
                np.add.at(nm, t0 + k_eff, -1)
                self.need.append(_runs(np.cumsum(nm)[:-1] > 0))
        finally:
            lib.bfCudaDebugFree(t)


def exchange_lists(all_have, all_need, ws_rows, arena, me: int):
    """For every stage t: (send_idx[q], recv_idx[p]) — workspace row indices this rank sends to q /
    receives from p after stage t, ascending (both sides derive the same lists).  A row produced at t
    in a ping-pong arena is wanted by the stage-(t+1) readers only; a dedicated row by any later
    reader.  Row ownership is disjoint, so a row src produced at t was not produced by dst."""
    world = len(all_have)
    nS = len(all_have[0])
    ded0 = arena[0] + arena[1]
    nd = ws_rows - ded0
    # dedicated rows needed at any stage > t, per rank (suffix OR), kept only for the dedicated tail
    later = [[None] * nS for _ in range(world)]
    for r in range(world):
        acc = np.zeros(nd, bool)
        for s in range(nS - 1, -1, -1):
            later[r][s] = acc.copy()
            if nd:
                acc |= _mask_from_runs(all_need[r][s], ws_rows)[ded0:]
    out = []
    for t in range(nS):
        send, recv = {}, {}
        have_me = _mask_from_runs(all_have[me][t], ws_rows)
        need_me = _mask_from_runs(all_need[me][t + 1], ws_rows) if t + 1 < nS else np.zeros(ws_rows, bool)
        need_me[ded0:] = later[me][t]
        for q in range(world):
            if q == me:
                continue
            want_q = _mask_from_runs(all_need[q][t + 1], ws_rows) if t + 1 < nS else np.zeros(ws_rows, bool)
            want_q[ded0:] = later[q][t]
            s_idx = np.nonzero(have_me & want_q)[0]
            r_idx = np.nonzero(_mask_from_runs(all_have[q][t], ws_rows) & need_me)[0]
            if len(s_idx):
                send[q] = s_idx
            if len(r_idx):
                recv[q] = r_idx
        out.append((send, recv))
    return out


# ---- runtime ----------------------------------------------------------------------------------------
class RowShardedOperator:
    """One rank's share of a block-row sharded operator.  `backend` runs stages and exposes the
    workspace: GpuBackend (bfCudaApplyStages) in the product, the table interpreter in the CPU tests."""

    def __init__(self, raw: RawOps, world: int, rank: int, backend_factory, num_queues: int = 148 * 12, group=None,
                 _pre=None, fractions=None):
        self.world, self.rank, self.group = world, rank, group
        if _pre is None:
            import torch.distributed as dist

            raw, self.owner, self.y_cuts = shard(raw, world, fractions)
            self.local = mask_for_rank(raw, self.owner, rank)
            st = StageTables(self.local, num_queues)
            gathered = [None] * world
            dist.all_gather_object(gathered, (st.have, st.need), group=group)
        else:  # in-process emulation of all ranks (tests on one GPU): tables computed by the caller
            raw, self.owner, self.y_cuts, self.local, st, gathered = _pre
        self.raw = raw
        dense = raw.ops["kind"] == capi.LEAF_DENSE
        self.local_entries = int((raw.ops["m"].astype(np.int64) * raw.ops["k"].astype(np.int64))[(self.owner == rank) & dense].sum())
        self.num_stages = st.num_stages
        self.lists = exchange_lists([g[0] for g in gathered], [g[1] for g in gathered], st.ws_rows, st.arena, rank)
        self.exchanged_rows = sum(len(v) for s, _ in self.lists for v in s.values())
        # every rank must enter the collective of stage t if any pair exchanges at t
        mine = [bool(s or r) for s, r in self.lists]
        if _pre is None:
            allf = [None] * world
            dist.all_gather_object(allf, mine, group=group)
            self.stage_exchanges = [any(f[t] for f in allf) for t in range(self.num_stages)]
        else:
            self.stage_exchanges = mine  # completed by build_all_local
        self.backend = backend_factory(self.local)
        self._idx_cache = {}

    @classmethod
    def build_all_local(cls, raw: RawOps, world: int, backend_factory, num_queues: int = 148 * 12, fractions=None):
        """All ranks' shares in ONE process, no process group: for emulating the sharded apply on a
        single device (emulate_apply)."""
        raw2, owner, y_cuts = shard(raw, world, fractions)
        locs = [mask_for_rank(raw2, owner, r) for r in range(world)]
        sts = [StageTables(l, num_queues) for l in locs]
        gathered = [(s.have, s.need) for s in sts]
        shares = [cls(raw, world, r, backend_factory, num_queues, None, _pre=(raw2, owner, y_cuts, locs[r], sts[r], gathered)) for r in range(world)]
        flags = [any(sh.stage_exchanges[t] for sh in shares) for t in range(shares[0].num_stages)]
        for sh in shares:
            sh.stage_exchanges = flags
        return shares

    @property
    def y_range(self):
        return self.y_cuts[self.rank], self.y_cuts[self.rank + 1]

    def apply(self, x, y):
        """y[own rows] = (A x)[own rows]; x is the full operand (replicated).  x, y: torch tensors."""
        import torch
        import torch.distributed as dist

        be = self.backend
        nrhs = x.shape[1]
        ws = be.begin(x, y)  # torch tensor view (ws_rows + pad, nrhs) of the workspace
        for t in range(self.num_stages):
            be.run_stage(t)
            send, recv = self.lists[t]
            if not self.stage_exchanges[t]:
                continue
            key = (t, ws.device)
            if key not in self._idx_cache:
                cat = lambda d: torch.as_tensor(np.concatenate([d[q] for q in sorted(d)]) if d else np.zeros(0, np.int64), device=ws.device)
                self._idx_cache[key] = (cat(send), cat(recv), [len(send.get(q, ())) * 1 for q in range(self.world)], [len(recv.get(q, ())) for q in range(self.world)])
            si, ri, ss, rs = self._idx_cache[key]
            sbuf = ws.index_select(0, si).contiguous()
            rbuf = torch.empty((int(ri.numel()), nrhs), dtype=ws.dtype, device=ws.device)
            # complex tensors travel as their real view (gloo has no complex all_to_all)
            sv = torch.view_as_real(sbuf).reshape(-1) if sbuf.is_complex() else sbuf.reshape(-1)
            rv = torch.view_as_real(rbuf).reshape(-1) if rbuf.is_complex() else rbuf.reshape(-1)
            e = (2 if sbuf.is_complex() else 1) * nrhs
            dist.all_to_all_single(rv, sv, [r * e for r in rs], [s * e for s in ss], group=self.group)
            if ri.numel():
                ws.index_copy_(0, ri, rbuf)
        be.end()
        return y

    def gather_rows(self, y, out=None):
        """Full result on every rank: one all_gather of the owned row ranges (padded to the longest when
        the ranges differ) through persistent staging buffers."""
        import torch
        import torch.distributed as dist

        if out is None:
            out = torch.empty_like(y)
        if self.world == 1:
            out.copy_(y)
            return out
        r0, r1 = self.y_range
        nrhs = y.shape[1]
        widths = [self.y_cuts[i + 1] - self.y_cuts[i] for i in range(self.world)]
        width = max(widths)
        real = lambda t: (torch.view_as_real(t) if t.is_complex() else t).reshape(-1)
        if min(widths) == width and y.is_contiguous() and out.is_contiguous():
            dist.all_gather_into_tensor(real(out), real(y[r0:r1]), group=self.group)  # no staging at all
            return out
        key = (y.device, y.dtype, nrhs)
        if not hasattr(self, "_gather_bufs"):
            self._gather_bufs = {}
        if key not in self._gather_bufs:
            self._gather_bufs[key] = (torch.zeros((width, nrhs), dtype=y.dtype, device=y.device),
                                      torch.empty((self.world * width, nrhs), dtype=y.dtype, device=y.device))
        pad_in, pad_out = self._gather_bufs[key]
        pad_in[: r1 - r0] = y[r0:r1]
        dist.all_gather_into_tensor(real(pad_out), real(pad_in), group=self.group)
        for i in range(self.world):
            a, b = self.y_cuts[i], self.y_cuts[i + 1]
            out[a:b] = pad_out[i * width: i * width + (b - a)]
        return out


class GpuBackend:
    """Stage runner over the C ABI (bfCudaApplyStages / bfCudaPlanGetWorkspace)."""

    def __init__(self, raw_local: RawOps, device: int = -1, synthetic_seed: int | None = None, helm2_recording=None):
        """helm2_recording: fill this rank's leaves with the real Helmholtz values (bfCudaPlanBuildHelm2 skips
        the REMOTE ones); the recording's leaf map must be for the same op list (shard() keeps a dense-leaf
        operator's ops as they are)."""
        from .plan import Plan

        self.plan = Plan(raw_local, device=device)
        self.build_stats = None
        if helm2_recording is not None:
            from . import helm2

            idx, off = origin_of(raw_local)  # leaves cut by shard() hold a row range of their recipe's matrix
            if len(helm2_recording.op_recipe) <= int(idx.max(initial=-1)):
                raise ValueError("the recording does not describe this op list")
            self.build_stats = helm2.build_plan(self.plan, helm2_recording.remapped(idx, off))
        elif synthetic_seed is not None:
            self.plan.fill_synthetic(synthetic_seed)
        self.lib = self.plan.lib

    def begin(self, x, y):
        import torch

        self.x, self.y = x, y
        nrhs = x.shape[1]
        p, ld = C.c_void_p(), C.c_int64()
        capi.check(self.lib.bfCudaPlanGetWorkspace(self.plan.handle, capi.OP_N, nrhs, C.byref(p), C.byref(ld)), "bfCudaPlanGetWorkspace")
        a0, a1, dd = C.c_int64(), C.c_int64(), C.c_int64()
        capi.check(self.lib.bfCudaPlanGetWorkspaceLayout(self.plan.handle, capi.OP_N, C.byref(a0), C.byref(a1), C.byref(dd)), "bfCudaPlanGetWorkspaceLayout")
        rows = a0.value + a1.value + dd.value + 2
        cplx = self.plan.dtype == capi.COMPLEX_F64

        class _Cai:
            __cuda_array_interface__ = {"shape": (rows, nrhs), "typestr": "<c16" if cplx else "<f8", "data": (p.value, False), "version": 3, "strides": None}

        self.stream = torch.cuda.current_stream().cuda_stream
        return torch.as_tensor(_Cai(), device=x.device)

    def run_stage(self, t):
        x, y = self.x, self.y
        s = C.c_void_p(self.stream if self.stream else 1)
        capi.check(
            self.lib.bfCudaApplyStages(self.plan.handle, capi.OP_N, C.c_void_p(x.data_ptr()), x.stride(0), x.shape[1],
                                       C.c_void_p(y.data_ptr()), y.stride(0), t, t + 1, s),
            "bfCudaApplyStages",
        )

    def end(self):
        pass

    def close(self):
        self.plan.close()


def emulate_apply(shares: List[RowShardedOperator], x, ys):
    """Run every rank's share in this process, stage by stage, moving the exchanged workspace rows with
    plain tensor copies instead of all_to_all_single.  Same kernels, same lists, one device."""
    import torch

    wss = [op.backend.begin(x, y) for op, y in zip(shares, ys)]
    for t in range(shares[0].num_stages):
        for op in shares:
            op.backend.run_stage(t)
        for src, op in enumerate(shares):
            for q, idx in op.lists[t][0].items():
                ridx = shares[q].lists[t][1][src]
                assert len(ridx) == len(idx)
                dev = wss[src].device
                wss[q].index_copy_(0, torch.as_tensor(ridx, device=dev), wss[src].index_select(0, torch.as_tensor(idx, device=dev)))
    for op in shares:
        op.backend.end()
    return ys
