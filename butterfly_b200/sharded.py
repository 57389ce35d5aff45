"""Host-side mirror of the multi-GPU C ABI (include/bf_cuda.h: bfCudaCommInit, bfCudaPlanShard,
bfCudaShardedApply*, and the per-rank bfCudaShardRank* used with one process per GPU).  ctypes only: ownership,
exchange set-up and the stage loop all live in csrc/bf_shard.cpp / bf_comm.cu."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .plan import Plan
from .rawops import RawOps


class Comm:
    """bfCudaCommInit: the GPUs of this process that share operators (an ordinal may repeat)."""

    def __init__(self, devices):
        self.lib = capi.load_library()
        devices = list(range(devices)) if isinstance(devices, int) else list(devices)
        arr = (C.c_int32 * len(devices))(*devices)
        self.handle = C.c_void_p()
        capi.check(self.lib.bfCudaCommInit(len(devices), arr, C.byref(self.handle)), "bfCudaCommInit")
        self.devices = devices

    def close(self):
        if getattr(self, "handle", None):
            self.lib.bfCudaCommDestroy(self.handle)
            self.handle = None

    __del__ = close


class ShardedPlan:
    """bfCudaPlanShard: one operator over the GPUs of a Comm, by block rows or by RHS columns."""

    def __init__(self, raw: RawOps, comm: Comm, mode: int = capi.SHARD_BLOCK_ROWS, fractions=None, build_transpose: int = 0):
        self.lib, self.comm, self.raw = capi.load_library(), comm, raw
        self.np_dtype = np.complex128 if raw.dtype == capi.COMPLEX_F64 else np.float64
        d = raw.desc(build_transpose=build_transpose)
        fr = None if fractions is None else (C.c_double * len(comm.devices))(*fractions)
        self.handle = C.c_void_p()
        capi.check(self.lib.bfCudaPlanShard(C.byref(d), comm.handle, mode, fr, C.byref(self.handle)), "bfCudaPlanShard")
        i = capi.ShardedInfo()
        capi.check(self.lib.bfCudaShardedPlanGetInfo(self.handle, C.byref(i)), "bfCudaShardedPlanGetInfo")
        self.info = i.as_dict()
        self.world = self.info["world"]

    def close(self):
        if getattr(self, "handle", None):
            self.lib.bfCudaShardedPlanDestroy(self.handle)
            self.handle = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def rank_plan(self, r: int) -> Plan:
        """Rank r's plan (borrowed): fill_synthetic / helm2 build / info."""
        return Plan.borrow(self.lib.bfCudaShardedPlanGetRankPlan(self.handle, r), self.raw.num_rows, self.raw.num_cols, self.raw.dtype)

    def apply_host(self, x: np.ndarray, op: int = capi.OP_N) -> np.ndarray:
        x = np.asarray(x)
        squeeze = x.ndim == 1
        x = np.ascontiguousarray(x[:, None] if squeeze else x, dtype=self.np_dtype)
        n_out = self.raw.num_rows if op == capi.OP_N else self.raw.num_cols
        y = np.empty((n_out, x.shape[1]), self.np_dtype)
        capi.check(self.lib.bfCudaShardedApplyHost(self.handle, op, x.ctypes.data, x.shape[1], x.shape[1], y.ctypes.data, x.shape[1]),
                   "bfCudaShardedApplyHost")
        return y[:, 0] if squeeze else y

    def apply_device(self, xs, ys):
        """Block rows: xs[r] / ys[r] = torch tensors on rank r's device (full operand / full-size result)."""
        px = (C.c_void_p * self.world)(*[t.data_ptr() for t in xs])
        py = (C.c_void_p * self.world)(*[t.data_ptr() for t in ys])
        capi.check(self.lib.bfCudaShardedApplyDevice(self.handle, px, py), "bfCudaShardedApplyDevice")

    def synchronize(self):
        capi.check(self.lib.bfCudaShardedSynchronize(self.handle), "bfCudaShardedSynchronize")

    def rank_stream(self, r: int) -> int:
        return int(self.lib.bfCudaShardedRankStream(self.handle, r) or 0)
