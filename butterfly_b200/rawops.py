"""Host-side container for a flattened operator: the raw-op list of include/bf_cuda.h.

`RawOps` owns numpy arrays (op table, virtual-buffer sizes, a packed value pool) and hands a
`BfCudaPlanDesc` to the C ABI.  It is how Python tests / benchmarks hold what the C flattener
(butterfly_b200/csrc/bf_flatten.c) produces from a reference `BfMat *`, how synthetic
factorizations are described, and what the golden fixtures under tests/golden/ store.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi


class RawOps:
    def __init__(self, dtype, num_rows, num_cols, buf_rows, ops, values=None):
        """ops: structured array (capi.RAWOP_DTYPE); if `values` is given (a float64 array; complex
        leaves are interleaved re, im), each dense/diag op's `val` field holds an offset IN DOUBLES
        into `values` (resolved to pointers by desc()); if values is None every val must be 0
        (value-less plan: the pool is filled by bfCudaPlanFillSynthetic)."""
        self.dtype = int(dtype)
        self.num_rows = int(num_rows)
        self.num_cols = int(num_cols)
        self.buf_rows = np.ascontiguousarray(buf_rows, dtype=np.int64)
        self.ops = np.ascontiguousarray(ops, dtype=capi.RAWOP_DTYPE)
        if values is not None:
            values = np.ascontiguousarray(values)
            if values.dtype == np.complex128:
                values = values.view(np.float64)
            assert values.dtype == np.float64 and values.ndim == 1
        self.values = values
        self._keep = None

    # ------------------------------------------------------------------------------------------
    @property
    def num_ops(self):
        return len(self.ops)

    @property
    def num_entries(self):
        d = self.ops[self.ops["kind"] == capi.LEAF_DENSE]
        return int((d["m"] * d["k"]).sum())

    def desc(self, device=-1, build_transpose=False, shard=None) -> capi.PlanDesc:
        """A ctypes BfCudaPlanDesc whose pointers stay valid while `self` is alive."""
        ops = self.ops.copy()
        if self.values is not None:
            base = self.values.ctypes.data
            has = ops["kind"] != capi.LEAF_IDENTITY
            ops["val"][has] = np.uint64(base) + ops["val"][has] * np.uint64(8)
            ops["val"][~has] = 0
        else:
            assert not ops["val"].any()
        d = capi.PlanDesc()
        d.abiVersion = capi.ABI_VERSION
        d.dtype = self.dtype
        d.numRows = self.num_rows
        d.numCols = self.num_cols
        d.numBufs = len(self.buf_rows)
        d.bufRows = self.buf_rows.ctypes.data_as(C.POINTER(C.c_int64))
        d.numOps = len(ops)
        d.ops = ops.ctypes.data_as(C.POINTER(capi.RawOp))
        d.device = device
        d.buildTranspose = int(build_transpose) | (capi.BUILD_KEEP_LEAVES if getattr(self, "keep_leaves", False) else 0)
        if shard is not None:  # (world, rank): one rank's share of a block-row sharded operator
            d.shardWorld, d.shardRank = int(shard[0]), int(shard[1])
        self._keep = ops  # the pointer table handed to C
        return d

    # ------------------------------------------------------------------------------------------
    @classmethod
    def from_desc(cls, desc: capi.PlanDesc) -> "RawOps":
        """Deep-copy a descriptor produced by the C flattener (values are gathered into one
        packed pool, each leaf re-packed with ld == stored column count)."""
        n = int(desc.numOps)
        buf = (capi.RawOp * n).from_address(C.addressof(desc.ops.contents)) if n else []
        ops = np.frombuffer(buf, dtype=capi.RAWOP_DTYPE).copy() if n else np.zeros(0, capi.RAWOP_DTYPE)
        buf_rows = np.ctypeslib.as_array(desc.bufRows, shape=(int(desc.numBufs),)).copy()
        cplx = desc.dtype == capi.COMPLEX_F64
        e = 2 if cplx else 1
        sizes = np.where(ops["kind"] == capi.LEAF_DENSE, ops["m"] * ops["k"] * e,
                         np.where(ops["kind"] == capi.LEAF_DIAG, np.minimum(ops["m"], ops["k"]), 0)).astype(np.int64) if n else np.zeros(0, np.int64)
        if n and not ops["val"][sizes > 0].any():
            # value-less descriptor (structure only): nothing to gather
            ops["val"] = 0
            return cls(desc.dtype, desc.numRows, desc.numCols, buf_rows, ops, None)
        offs = np.concatenate([[0], np.cumsum(sizes)])
        pool = np.zeros(int(offs[-1]), np.float64)
        for i in range(n):
            o = ops[i]
            if sizes[i] == 0 or int(o["val"]) == 0:
                ops["val"][i] = 0
                continue
            if o["kind"] == capi.LEAF_DENSE:
                acts_t = bool(o["flags"] & (capi.LEAF_TRANS | (capi.LEAF_CONJ if cplx else 0)))
                sr, sc = (int(o["k"]), int(o["m"])) if acts_t else (int(o["m"]), int(o["k"]))
                ld = int(o["ld"])
                raw = np.ctypeslib.as_array(
                    C.cast(int(o["val"]), C.POINTER(C.c_double)), shape=(((sr - 1) * ld + sc) * e,)
                )
                blk = np.lib.stride_tricks.as_strided(raw, (sr, sc * e), (ld * e * 8, 8))
                pool[offs[i] : offs[i + 1]] = blk.reshape(-1)
                ops["ld"][i] = sc
            else:  # diag: real values
                raw = np.ctypeslib.as_array(
                    C.cast(int(o["val"]), C.POINTER(C.c_double)), shape=(int(sizes[i]),)
                )
                pool[offs[i] : offs[i + 1]] = raw
            ops["val"][i] = offs[i]
        return cls(desc.dtype, desc.numRows, desc.numCols, buf_rows, ops, pool)

    # ------------------------------------------------------------------------------------------
    def save(self, path):
        np.savez_compressed(
            path,
            dtype=self.dtype,
            num_rows=self.num_rows,
            num_cols=self.num_cols,
            buf_rows=self.buf_rows,
            ops=self.ops,
            values=self.values if self.values is not None else np.zeros(0),
            has_values=self.values is not None,
        )

    @classmethod
    def load(cls, path) -> "RawOps":
        z = np.load(path, allow_pickle=False)
        vals = z["values"] if bool(z["has_values"]) else None
        return cls(int(z["dtype"]), int(z["num_rows"]), int(z["num_cols"]), z["buf_rows"],
                   z["ops"].astype(capi.RAWOP_DTYPE), vals)

    # ------------------------------------------------------------------------------------------
    def fingerprint(self) -> np.ndarray:
        """All integer content (bit-exact comparison of index handling)."""
        o = self.ops
        cols = [o[f].astype(np.int64) for f in ("kind", "flags", "inBuf", "inOff", "outBuf", "outOff", "m", "k")]
        return np.stack(cols, axis=1) if len(o) else np.zeros((0, 8), np.int64)
