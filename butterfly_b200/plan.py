"""Host-side handle on a device-resident butterfly plan (BfCudaPlan), through the C ABI only.

Mirrors what a C caller of include/bf_cuda.h does: create from a raw-op descriptor, apply to host
buffers (bfCudaApplyHost — the call the reference-facing bfMatMul override makes) or to device
buffers (bfCudaApplyDevice — the timed entry point), destroy.  torch is used only to own device
memory / streams on the caller's side.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .rawops import RawOps


class Plan:
    def __init__(self, raw: RawOps, device: int = -1, build_transpose: int = 0):
        self.lib = capi.load_library()
        self.raw = raw
        self.dtype = raw.dtype
        self.np_dtype = np.complex128 if raw.dtype == capi.COMPLEX_F64 else np.float64
        desc = raw.desc(device=device, build_transpose=build_transpose)
        h = C.c_void_p()
        capi.check(self.lib.bfCudaPlanCreate(C.byref(desc), C.byref(h)), "bfCudaPlanCreate")
        self.handle = h
        self._info = None

    @classmethod
    def borrow(cls, handle, num_rows: int, num_cols: int, dtype: int = capi.COMPLEX_F64) -> "Plan":
        """A handle on a plan somebody else owns (bfCudaMatGetPlan of an attached reference object):
        same methods, never destroyed from here."""
        self = cls.__new__(cls)
        self.lib = capi.load_library()
        self.raw = type("Shape", (), {"num_rows": int(num_rows), "num_cols": int(num_cols), "dtype": dtype})()
        self.dtype = dtype
        self.np_dtype = np.complex128 if dtype == capi.COMPLEX_F64 else np.float64
        self.handle = C.c_void_p(handle) if not isinstance(handle, C.c_void_p) else handle
        self._info = None
        self._borrowed = True
        return self

    def save(self, path):
        """bfCudaPlanSave: tables + the device-resident factor pool, every built direction."""
        capi.check(self.lib.bfCudaPlanSave(self.handle, str(path).encode()), "bfCudaPlanSave")

    @classmethod
    def load(cls, path, device: int = -1) -> "Plan":
        """bfCudaPlanLoad: a plan that applies bit-identically to the saved one."""
        lib = capi.load_library()
        h = C.c_void_p()
        capi.check(lib.bfCudaPlanLoad(str(path).encode(), device, C.byref(h)), "bfCudaPlanLoad")
        i = capi.PlanInfo()
        capi.check(lib.bfCudaPlanGetInfo(h, C.byref(i)), "bfCudaPlanGetInfo")
        self = cls.borrow(h, i.numRows, i.numCols, i.dtype)
        self._borrowed = False
        return self

    # ------------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "handle", None):
            if not getattr(self, "_borrowed", False):
                self.lib.bfCudaPlanDestroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ------------------------------------------------------------------------------------------
    @property
    def info(self) -> dict:
        if self._info is None:
            i = capi.PlanInfo()
            capi.check(self.lib.bfCudaPlanGetInfo(self.handle, C.byref(i)), "bfCudaPlanGetInfo")
            self._info = i.as_dict()
        return self._info

    def rows_in(self, op):
        return self.raw.num_cols if op == capi.OP_N else self.raw.num_rows

    def rows_out(self, op):
        return self.raw.num_rows if op == capi.OP_N else self.raw.num_cols

    def set_perms(self, in_perm=None, out_perm=None):
        def ptr(p):
            if p is None:
                return None, None
            a = np.ascontiguousarray(p, dtype=np.int64)
            return a, a.ctypes.data_as(C.POINTER(C.c_int64))

        ia, ip = ptr(in_perm)
        oa, op = ptr(out_perm)
        capi.check(self.lib.bfCudaPlanSetPerms(self.handle, ip, op), "bfCudaPlanSetPerms")

    def fill_synthetic(self, seed: int):
        capi.check(self.lib.bfCudaPlanFillSynthetic(self.handle, C.c_uint64(seed)), "bfCudaPlanFillSynthetic")

    def set_fused(self, on: bool):
        """1-RHS applies as one cooperative launch (default) or one launch per stage."""
        capi.check(self.lib.bfCudaPlanSetFused(self.handle, int(on)), "bfCudaPlanSetFused")

    def is_fused(self, op: int = capi.OP_N) -> bool:
        return bool(self.lib.bfCudaPlanGetFused(self.handle, op))

    def set_profiling(self, on: bool):
        capi.check(self.lib.bfCudaPlanSetProfiling(self.handle, int(on)), "bfCudaPlanSetProfiling")

    def stage_times_ms(self) -> np.ndarray:
        n = C.c_int64()
        capi.check(self.lib.bfCudaPlanGetStageTimes(self.handle, None, 0, C.byref(n)), "bfCudaPlanGetStageTimes")
        ms = np.zeros(n.value, np.float32)
        capi.check(
            self.lib.bfCudaPlanGetStageTimes(self.handle, ms.ctypes.data_as(C.POINTER(C.c_float)), n.value, C.byref(n)),
            "bfCudaPlanGetStageTimes",
        )
        return ms

    def stage_entries(self, op=capi.OP_N) -> np.ndarray:
        n = C.c_int64()
        capi.check(self.lib.bfCudaPlanGetStageEntries(self.handle, op, None, 0, C.byref(n)), "bfCudaPlanGetStageEntries")
        e = np.zeros(n.value, np.int64)
        capi.check(
            self.lib.bfCudaPlanGetStageEntries(self.handle, op, e.ctypes.data_as(C.POINTER(C.c_int64)), n.value, C.byref(n)),
            "bfCudaPlanGetStageEntries",
        )
        return e

    @property
    def last_launch_count(self) -> int:
        return int(self.lib.bfCudaPlanLastLaunchCount(self.handle))

    # ------------------------------------------------------------------------------------------
    def apply_host(self, x: np.ndarray, op: int = capi.OP_N, out: np.ndarray | None = None) -> np.ndarray:
        """y = op(A) x on host buffers (H2D, apply, D2H inside the call)."""
        x = np.asarray(x)
        squeeze = x.ndim == 1
        if squeeze:
            x = x[:, None]
        if x.dtype != self.np_dtype or x.strides[1] != x.itemsize or x.strides[0] % x.itemsize or x.strides[0] < x.shape[1] * x.itemsize:
            x = np.ascontiguousarray(x, dtype=self.np_dtype)
        n, nrhs = x.shape
        if n != self.rows_in(op):
            raise ValueError(f"operand has {n} rows, operator expects {self.rows_in(op)}")
        y = out if out is not None else np.empty((self.rows_out(op), nrhs), self.np_dtype)
        ldx = x.strides[0] // x.itemsize if n > 1 else max(nrhs, 1)
        ldy = y.strides[0] // y.itemsize if y.shape[0] > 1 else max(nrhs, 1)
        capi.check(
            self.lib.bfCudaApplyHost(self.handle, op, x.ctypes.data, ldx, nrhs, y.ctypes.data, ldy),
            "bfCudaApplyHost",
        )
        return y[:, 0] if squeeze and out is None else y

    def solve_gmres(self, b: np.ndarray, x0: np.ndarray | None = None, tol: float = 1e-12, max_iter: int = 50):
        """bfCudaSolveGMRES: the algorithm of bfSolveGMRES (src/linalg.c:47) with the Krylov basis
        resident in HBM.  Returns (x, num_iter, rel_residual, num_applies)."""
        b = np.ascontiguousarray(b, dtype=np.complex128).reshape(-1)
        x0a = None if x0 is None else np.ascontiguousarray(x0, dtype=np.complex128).reshape(-1)
        x = np.empty_like(b)
        it, res, na = C.c_int64(), C.c_double(), C.c_int64()
        capi.check(
            self.lib.bfCudaSolveGMRES(self.handle, b.ctypes.data, None if x0a is None else x0a.ctypes.data, float(tol), int(max_iter),
                                      x.ctypes.data, C.byref(it), C.byref(res), C.byref(na)),
            "bfCudaSolveGMRES",
        )
        return x, it.value, res.value, na.value

    def apply_device(self, x, y, op: int = capi.OP_N, stream=None):
        """y = op(A) x on torch CUDA tensors (row-major n x nrhs, last stride 1).  Asynchronous."""
        assert x.is_cuda and y.is_cuda and x.stride(-1) == 1 and y.stride(-1) == 1
        nrhs = x.shape[1] if x.dim() == 2 else 1
        ldx = x.stride(0) if x.dim() == 2 else 1
        ldy = y.stride(0) if y.dim() == 2 else 1
        if x.shape[0] != self.rows_in(op) or y.shape[0] != self.rows_out(op):
            raise ValueError("shape mismatch")
        # NULL means "the plan's own stream" in the C ABI; torch's default stream has handle 0,
        # which is the legacy default stream = cudaStreamLegacy (0x1) as an explicit handle
        s = None if stream is None else C.c_void_p(stream if stream else 1)
        capi.check(
            self.lib.bfCudaApplyDevice(self.handle, op, C.c_void_p(x.data_ptr()), ldx, nrhs, C.c_void_p(y.data_ptr()), ldy, s),
            "bfCudaApplyDevice",
        )
        return y
