"""ctypes wrapper of oracle/libplan_oracle.so — the plain-C CPU restatement ("port") of the
reference's apply over a flattened operator.  TEST INFRASTRUCTURE, NOT PRODUCT: only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from butterfly_b200 import capi
from butterfly_b200.rawops import RawOps

HERE = Path(__file__).resolve().parent
_lib = None


def lib():
    global _lib
    if _lib is None:
        so = HERE / "libplan_oracle.so"
        if not so.exists():
            subprocess.run(["make", "-C", str(HERE), "port"], check=True, stdout=subprocess.DEVNULL)
        l = C.CDLL(str(so))
        l.planOracleApply.argtypes = [C.POINTER(capi.PlanDesc), C.c_int32, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int64]
        l.planOracleApply.restype = C.c_int
        l.planOracleNumEntries.argtypes = [C.POINTER(capi.PlanDesc)]
        l.planOracleNumEntries.restype = C.c_int64
        l.planOraclePermuteRows.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
        l.planOraclePermuteRows.restype = None
        _lib = l
    return _lib


def apply(raw: RawOps, x: np.ndarray, op: int = capi.OP_N) -> np.ndarray:
    """y = op(A) x by the scalar-C oracle, leaves summed in the reference's traversal order."""
    assert raw.values is not None, "materialize() synthetic raw ops first"
    dt = np.complex128 if raw.dtype == capi.COMPLEX_F64 else np.float64
    x = np.ascontiguousarray(x, dtype=dt)
    squeeze = x.ndim == 1
    if squeeze:
        x = x[:, None]
    nrhs = x.shape[1]
    rows_out = raw.num_rows if op == capi.OP_N else raw.num_cols
    y = np.zeros((rows_out, nrhs), dt)
    d = raw.desc()
    rc = lib().planOracleApply(C.byref(d), op, x.ctypes.data, max(nrhs, 1), nrhs, y.ctypes.data, max(nrhs, 1))
    if rc != 0:
        raise RuntimeError(f"planOracleApply failed: {rc}")
    return y[:, 0] if squeeze else y


def permute_rows(perm, x: np.ndarray) -> np.ndarray:
    """out[perm[i], :] = x[i, :] (bfMatDenseComplexPermuteRows / bfVecRealPermute semantics)."""
    x2 = np.ascontiguousarray(x if x.ndim == 2 else x[:, None])
    out = np.empty_like(x2)
    p = np.ascontiguousarray(perm, dtype=np.int64)
    ed = 2 if np.iscomplexobj(x2) else 1
    lib().planOraclePermuteRows(p.ctypes.data, len(p), x2.shape[1], ed, x2.ctypes.data, x2.shape[1], out.ctypes.data, x2.shape[1])
    return out if x.ndim == 2 else out[:, 0]


# ---- synthetic values: numpy restatement of syntheticValue() in butterfly_b200/csrc/bf_plan.hpp ----
def _hash_values(seed: int, raw_index: int, linear: np.ndarray, part: int, scale: float) -> np.ndarray:
    M = np.uint64
    with np.errstate(over="ignore"):
        z = M(seed) ^ (M(0x9E3779B97F4A7C15) * M(raw_index + 1))
        z = z + M(0xBF58476D1CE4E5B9) * (M(2) * linear.astype(np.uint64) + M(part + 1))
        z = (z ^ (z >> M(30))) * M(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> M(27))) * M(0x94D049BB133111EB)
        z = z ^ (z >> M(31))
    u = (z >> M(11)).astype(np.int64).astype(np.float64) * (1.0 / 4503599627370496.0) - 1.0
    return u * scale


def synthetic_scale(k: int) -> float:
    e = int(k).bit_length() - 1
    return 1.0 / float(1 << ((e + 1) // 2))


def materialize(raw: RawOps, seed: int) -> RawOps:
    """Host copy of what bfCudaPlanFillSynthetic(seed) puts on the device for a value-less plan."""
    cplx = raw.dtype == capi.COMPLEX_F64
    e = 2 if cplx else 1
    ops = raw.ops.copy()
    sizes = np.where(ops["kind"] == capi.LEAF_DENSE, ops["m"] * ops["k"] * e,
                     np.where(ops["kind"] == capi.LEAF_DIAG, np.minimum(ops["m"], ops["k"]), 0)).astype(np.int64)
    offs = np.concatenate([[0], np.cumsum(sizes)])
    pool = np.zeros(int(offs[-1]), np.float64)
    for i in range(len(ops)):
        n = int(sizes[i])
        if n == 0:
            ops["val"][i] = 0
            continue
        if ops["kind"][i] == capi.LEAF_DIAG:
            pool[offs[i]:offs[i + 1]] = 1.0 + 0.5 * _hash_values(seed, i, np.arange(n, dtype=np.uint64), 0, 1.0)
        else:
            lin = np.arange(n // e, dtype=np.uint64)
            sc = synthetic_scale(int(ops["k"][i]))
            seg = pool[offs[i]:offs[i + 1]]
            if cplx:
                seg[0::2] = _hash_values(seed, i, lin, 0, sc)
                seg[1::2] = _hash_values(seed, i, lin, 1, sc)
            else:
                seg[:] = _hash_values(seed, i, lin, 0, sc)
            mask = (capi.LEAF_TRANS | capi.LEAF_CONJ) if cplx else capi.LEAF_TRANS
            st = bool(ops["flags"][i] & mask)
            ops["ld"][i] = ops["m"][i] if st else ops["k"][i]
        ops["val"][i] = offs[i]
    return RawOps(raw.dtype, raw.num_rows, raw.num_cols, raw.buf_rows, ops, pool)
