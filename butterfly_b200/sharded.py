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


class ShardRankOperator:
    """ONE rank of a block-row sharded operator, for programs that run one process per GPU (bench.py under
    torchrun).  Ownership comes from bfCudaShardDescsCreate (the same on every rank), the rank's plan, workspace
    and flags from bfCudaShardRankCreate; the ranks swap BfCudaShardHandle structs once (`exchange`: a callable
    that all-gathers a bytes object, e.g. over torch.distributed) and connect through CUDA IPC.  After that an
    apply is `numStages` kernel launches on this rank's stream and nothing else: rows other ranks read are stored
    into their workspaces by the kernels, no collective runs.

    Same duck type as rowshard.RowShardedOperator (world, rank, y_cuts, y_range, group, apply, gather_rows), so
    butterfly_b200/gmres.py runs over it unchanged."""

    def __init__(self, raw: RawOps, world: int, rank: int, device: int, exchange, fractions=None, group=None,
                 synthetic_seed: int | None = None, helm2_recording=None):
        lib = self.lib = capi.load_library()
        self.world, self.rank, self.group = world, rank, group
        d = raw.desc()
        fr = None if fractions is None else (C.c_double * world)(*fractions)
        sd = C.c_void_p()
        capi.check(lib.bfCudaShardDescsCreate(C.byref(d), world, fr, C.byref(sd)), "bfCudaShardDescsCreate")
        try:
            n = C.c_int64()
            capi.check(lib.bfCudaShardDescsGetLayout(sd, C.byref(n), None, None, None, None), "bfCudaShardDescsGetLayout")
            cuts = (C.c_int64 * (world + 1))()
            self.owner = np.zeros(n.value, np.int32)
            self.origin_op, self.origin_row = np.zeros(n.value, np.int64), np.zeros(n.value, np.int64)
            p32, p64 = C.POINTER(C.c_int32), C.POINTER(C.c_int64)
            capi.check(lib.bfCudaShardDescsGetLayout(sd, None, cuts, self.owner.ctypes.data_as(p32), self.origin_op.ctypes.data_as(p64),
                                                     self.origin_row.ctypes.data_as(p64)), "bfCudaShardDescsGetLayout")
            self.y_cuts = [int(c) for c in cuts]
            rd = capi.PlanDesc.from_buffer_copy(lib.bfCudaShardDescsGet(sd, rank).contents)
            rd.device = device
            # the sharded op list without the ownership marks: the UNSHARDED operator with the same leaves (and the
            # same raw indices, which is what synthetic values hash on) — what a verification plan is built from
            self.raw = RawOps.from_desc(rd)
            self.raw.ops["flags"] &= np.uint32(~(capi.LEAF_REMOTE | (0xFF << 8)) & 0xFFFFFFFF)
            self.handle = C.c_void_p()
            capi.check(lib.bfCudaShardRankCreate(C.byref(rd), C.byref(self.handle)), "bfCudaShardRankCreate")
        finally:
            lib.bfCudaShardDescsFree(sd)
        self.plan = Plan.borrow(lib.bfCudaShardRankGetPlan(self.handle), raw.num_rows, raw.num_cols, raw.dtype)
        self.build_stats = None
        if helm2_recording is not None:
            from . import helm2

            self.build_stats = helm2.build_plan(self.plan, helm2_recording.remapped(self.origin_op, self.origin_row))
        elif synthetic_seed is not None:
            self.plan.fill_synthetic(synthetic_seed)
        info = self.plan.info
        self.local_entries = int(info["numEntries"])
        self.num_stages = int(info["numStagesN"])
        self.exchanged_rows = int(lib.bfCudaShardRankSentRows(self.handle))
        mine = capi.ShardHandle()
        capi.check(lib.bfCudaShardRankGetHandle(self.handle, C.byref(mine)), "bfCudaShardRankGetHandle")
        blobs = exchange(bytes(mine))  # [world] bytes, in rank order
        allh = (capi.ShardHandle * world)(*[capi.ShardHandle.from_buffer_copy(b) for b in blobs])
        capi.check(lib.bfCudaShardRankConnect(self.handle, allh), "bfCudaShardRankConnect")

    @property
    def y_range(self):
        return self.y_cuts[self.rank], self.y_cuts[self.rank + 1]

    def apply(self, x, y, stream=None):
        """y[own rows] = (A x)[own rows]; x: the full operand (n x 1 torch tensor on this rank's device)."""
        import torch

        assert x.shape[1] == 1 and x.is_contiguous() and y.is_contiguous()
        s = torch.cuda.current_stream().cuda_stream if stream is None else stream
        capi.check(self.lib.bfCudaShardRankApplyDevice(self.handle, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), C.c_void_p(s or 1)),
                   "bfCudaShardRankApplyDevice")
        return y

    def gather_rows(self, y, out=None):
        from .rowshard import RowShardedOperator

        return RowShardedOperator.gather_rows(self, y, out)

    def close(self):
        if getattr(self, "handle", None):
            self.lib.bfCudaShardRankDestroy(self.handle)
            self.handle = None
