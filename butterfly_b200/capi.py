"""ctypes binding of the C ABI in include/bf_cuda.h (libbfcuda.so).

This is the only way Python code in this repo reaches the product: tests, bench.py and
__graft_entry__ all call the same `extern "C"` entry points a C host program would.  There is no
fallback of any kind: if the shared library is missing, importing a symbol from it raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
REPO_ROOT = PKG_DIR.parent
LIB_DIR = PKG_DIR / "lib"
LIB_PATH = LIB_DIR / "libbfcuda.so"
ATTACH_LIB_PATH = LIB_DIR / "libbfcuda_attach.so"

ABI_VERSION = 2

# enums (include/bf_cuda.h)
OK = 0
REAL_F64, COMPLEX_F64 = 0, 1
OP_N, OP_T, OP_H = 0, 1, 2
LEAF_DENSE, LEAF_IDENTITY, LEAF_DIAG = 0, 1, 2
BUILD_KEEP_LEAVES = 4  # BF_CUDA_BUILD_KEEP_LEAVES: PlanDesc.buildTranspose bit 2
LEAF_TRANS, LEAF_CONJ, LEAF_REMOTE = 2, 4, 8
BUF_X, BUF_Y, BUF_FIRST_TEMP = 0, 1, 2


class RawOp(C.Structure):
    """struct BfCudaRawOp"""

    _fields_ = [
        ("kind", C.c_int32),
        ("flags", C.c_uint32),
        ("inBuf", C.c_int64),
        ("inOff", C.c_int64),
        ("outBuf", C.c_int64),
        ("outOff", C.c_int64),
        ("m", C.c_int64),
        ("k", C.c_int64),
        ("ld", C.c_int64),
        ("val", C.c_void_p),
    ]


RAWOP_DTYPE = np.dtype(
    [
        ("kind", "<i4"),
        ("flags", "<u4"),
        ("inBuf", "<i8"),
        ("inOff", "<i8"),
        ("outBuf", "<i8"),
        ("outOff", "<i8"),
        ("m", "<i8"),
        ("k", "<i8"),
        ("ld", "<i8"),
        ("val", "<u8"),
    ],
    align=True,
)
assert RAWOP_DTYPE.itemsize == C.sizeof(RawOp) == 72


class PlanDesc(C.Structure):
    """struct BfCudaPlanDesc"""

    _fields_ = [
        ("abiVersion", C.c_int32),
        ("dtype", C.c_int32),
        ("numRows", C.c_int64),
        ("numCols", C.c_int64),
        ("numBufs", C.c_int64),
        ("bufRows", C.POINTER(C.c_int64)),
        ("numOps", C.c_int64),
        ("ops", C.POINTER(RawOp)),
        ("device", C.c_int32),
        ("buildTranspose", C.c_int32),
        ("shardWorld", C.c_int32),
        ("shardRank", C.c_int32),
    ]


class PlanInfo(C.Structure):
    """struct BfCudaPlanInfo"""

    _fields_ = [
        ("dtype", C.c_int32),
        ("device", C.c_int32),
        ("numRows", C.c_int64),
        ("numCols", C.c_int64),
        ("numOps", C.c_int64),
        ("numEntries", C.c_int64),
        ("numStagesN", C.c_int64),
        ("numStagesT", C.c_int64),
        ("numItemsN", C.c_int64),
        ("numTilesN", C.c_int64),
        ("poolBytes", C.c_int64),
        ("tableBytes", C.c_int64),
        ("workspaceRowsN", C.c_int64),
        ("workspaceRowsT", C.c_int64),
        ("maxTileK", C.c_int64),
        ("maxItemRows", C.c_int64),
    ]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class DebugTables(C.Structure):
    """struct BfCudaDebugTables (include/bf_cuda_debug.h)"""

    _fields_ = [
        ("dtype", C.c_int32),
        ("op", C.c_int32),
        ("numStages", C.c_int32),
        ("numQueues", C.c_int32),
        ("rowsIn", C.c_int64),
        ("rowsOut", C.c_int64),
        ("wsRows", C.c_int64),
        ("numItems", C.c_int64),
        ("numTiles", C.c_int64),
        ("stageItemBegin", C.POINTER(C.c_int64)),
        ("stageTileBegin", C.POINTER(C.c_int64)),
        ("items", C.c_void_p),
        ("tiles", C.c_void_p),
        ("tilePool16", C.POINTER(C.c_uint64)),
        ("qItemBegin", C.POINTER(C.c_uint32)),
        ("qTileBegin", C.POINTER(C.c_uint32)),
        ("qPool16", C.POINTER(C.c_uint64)),
        ("tileOp", C.POINTER(C.c_int64)),
        ("tileRow0", C.POINTER(C.c_int32)),
        ("tileCol0", C.POINTER(C.c_int32)),
        ("pool16", C.c_uint64),
        ("pool", C.c_void_p),
        ("wsArena0Rows", C.c_int64),
        ("wsArena1Rows", C.c_int64),
        ("owner", C.c_void_p),
        ("fusedEnabled", C.c_int32),
        ("fusedNumCounters", C.c_uint32),
        ("fusedWsRows", C.c_int64),
        ("fusedNumItems", C.c_int64),
        ("fusedNumTiles", C.c_int64),
        ("fusedItems", C.c_void_p),
        ("fusedItemDeps", C.c_void_p),
        ("fusedTiles", C.c_void_p),
        ("fusedTileDeps", C.c_void_p),
        ("fusedQItemBegin", C.POINTER(C.c_uint32)),
        ("fusedQTileBegin", C.POINTER(C.c_uint32)),
        ("itemPeers", C.POINTER(C.c_uint32)),
        ("stageRecvMask", C.POINTER(C.c_uint32)),
        ("stageSendMask", C.POINTER(C.c_uint32)),
    ]


class ShardHandle(C.Structure):
    """struct BfCudaShardHandle"""

    _fields_ = [
        ("pid", C.c_int64), ("device", C.c_int32), ("rank", C.c_int32), ("wsBytes", C.c_int64),
        ("wsPtr", C.c_void_p), ("flagsPtr", C.c_void_p), ("hasIpc", C.c_int32), ("reserved", C.c_int32),
        ("wsIpc", C.c_ubyte * 64), ("flagsIpc", C.c_ubyte * 64),
    ]


class ShardedInfo(C.Structure):
    """struct BfCudaShardedInfo"""

    _fields_ = [
        ("world", C.c_int32), ("mode", C.c_int32), ("numRows", C.c_int64), ("numCols", C.c_int64),
        ("numEntries", C.c_int64), ("maxRankEntries", C.c_int64), ("poolBytesMax", C.c_int64),
        ("sentRowsPerApply", C.c_int64), ("numStages", C.c_int64), ("yCuts", C.c_int64 * 17),
    ]

    def as_dict(self):
        d = {n: getattr(self, n) for n, _ in self._fields_ if n != "yCuts"}
        d["yCuts"] = list(self.yCuts)[: self.world + 1]
        return d


SHARD_RHS_COLS, SHARD_BLOCK_ROWS = 0, 1


ITEM_DTYPE = np.dtype([("outAddr", "<u4"), ("tileBegin", "<u4"), ("rows", "<u2"), ("flags", "<u2"), ("numTiles", "<u4")])
TILE_DTYPE = np.dtype([("inAddr", "<u4"), ("kc", "<u2"), ("rows", "<u2")])
KC_IDENTITY, KC_DIAG, KC_JUMP = 0xFFFF, 0xFFFE, 0xFFFD
ITEM_ACCUMULATE, ITEM_JUMP = 1, 2
NO_COUNTER = 0xFFFFFFFF
ITEM_DEP_DTYPE = np.dtype([("sig", "<u4"), ("waitTarget", "<u4")])
TILE_DEP_DTYPE = np.dtype([("idx", "<u4"), ("total", "<u4")])
SPACE_X, SPACE_Y, SPACE_W = 0, 1, 2
ROW_MASK = (1 << 30) - 1

# every symbol include/bf_cuda.h declares; tests check that the .so exports all of them
EXPORTED_SYMBOLS = [
    "bfCudaPlanCreate",
    "bfCudaPlanDestroy",
    "bfCudaPlanGetInfo",
    "bfCudaPlanFillSynthetic",
    "bfCudaSyntheticValue",
    "bfCudaSyntheticDiagValue",
    "bfCudaPlanSetPerms",
    "bfCudaApplyDevice",
    "bfCudaApplyHost",
    "bfCudaPlanLastLaunchCount",
    "bfCudaMeasureFp64Peak",
    "bfCudaMeasureFp64PeakSustained",
    "bfCudaPlanGetStream",
    "bfCudaPlanSetFused",
    "bfCudaPlanGetFused",
    "bfCudaApplyStages",
    "bfCudaPlanGetWorkspace",
    "bfCudaPlanGetWorkspaceLayout",
    "bfCudaSolveGMRES",
    "bfCudaSolveGMRESPrecond",
    "bfCudaGmresScratchBytes",
    "bfCudaGmresProject",
    "bfCudaGmresSubtract",
    "bfCudaGmresScale",
    "bfCudaGmresProjectStatic",
    "bfCudaShardDescsCreate",
    "bfCudaShardDescsGet",
    "bfCudaShardDescsGetLayout",
    "bfCudaShardDescsFree",
    "bfCudaShardRankCreate",
    "bfCudaShardRankDestroy",
    "bfCudaShardRankGetPlan",
    "bfCudaShardRankSentRows",
    "bfCudaShardRankGetHandle",
    "bfCudaShardRankConnect",
    "bfCudaShardRankApplyDevice",
    "bfCudaShardRankApplyStages",
    "bfCudaCommInit",
    "bfCudaCommDestroy",
    "bfCudaCommSize",
    "bfCudaPlanShard",
    "bfCudaShardedPlanDestroy",
    "bfCudaShardedPlanGetInfo",
    "bfCudaShardedPlanGetRankPlan",
    "bfCudaShardedApplyHost",
    "bfCudaShardedApplyDevice",
    "bfCudaShardedSynchronize",
    "bfCudaShardedRankStream",
    "bfCudaPlanSetProfiling",
    "bfCudaPlanGetStageTimes",
    "bfCudaPlanGetStageEntries",
    "bfCudaPlanBuildHelm2",
    "bfCudaHelm2Evaluate",
    "bfCudaPlanSave",
    "bfCudaPlanLoad",
    "bfCudaShardRankLoad",
    "bfCudaRawOpsSave",
    "bfCudaRawOpsLoad",
    "bfCudaRawOpsFree",
    "bfCudaStatusString",
    "bfCudaLastErrorMessage",
    "bfCudaAbiVersion",
]
DEBUG_EXPORTED_SYMBOLS = ["bfCudaDebugCompile", "bfCudaDebugFree"]
ATTACH_EXPORTED_SYMBOLS = [
    "bfCudaFlatten",
    "bfCudaFlattenEx",
    "bfCudaFlattenFree",
    "bfCudaMatAttach",
    "bfCudaMatDetach",
    "bfCudaMatGetPlan",
    "bfCudaMatSetPerms",
    "bfCudaSolveGMRESMat",
    "bfCudaMatAttachHelm2",
    "bfCudaMatAttachSharded",
    "bfCudaMatGetShardedPlan",
]
# libbfcuda_helm2.so (csrc/bf_helm2_record.c): the recorder of the reference's Helmholtz value builders
HELM2_EXPORTED_SYMBOLS = [
    "bfCudaHelm2RecordBegin",
    "bfCudaHelm2RecordEnd",
    "bfCudaHelm2RecordingFree",
    "bfCudaHelm2RecordingSetTruncationScale",
    "bfCudaHelm2Resolve",
    "bfHelm2GetKernelMatrix",
    "bfHelm2GetReexpansionMatrix",
]

_lib = None


class BfCudaError(RuntimeError):
    def __init__(self, status, where, detail=""):
        self.status = status
        super().__init__(f"{where}: status {status} ({detail})")


def load_library(path: os.PathLike | None = None) -> C.CDLL:
    """dlopen libbfcuda.so and declare prototypes.  Raises if the library is absent: the product
    has no other path."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = Path(path) if path else LIB_PATH
    if not p.exists():
        raise FileNotFoundError(
            f"{p} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)"
        )
    lib = C.CDLL(str(p), mode=C.RTLD_GLOBAL)
    vp, i32, i64, u64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64
    lib.bfCudaPlanCreate.argtypes = [C.POINTER(PlanDesc), C.POINTER(vp)]
    lib.bfCudaPlanCreate.restype = i32
    lib.bfCudaPlanDestroy.argtypes = [vp]
    lib.bfCudaPlanDestroy.restype = i32
    lib.bfCudaPlanGetInfo.argtypes = [vp, C.POINTER(PlanInfo)]
    lib.bfCudaPlanGetInfo.restype = i32
    lib.bfCudaPlanFillSynthetic.argtypes = [vp, u64]
    lib.bfCudaPlanFillSynthetic.restype = i32
    lib.bfCudaSyntheticValue.argtypes = [u64, i64, i64, i32, i64]
    lib.bfCudaSyntheticValue.restype = C.c_double
    lib.bfCudaSyntheticDiagValue.argtypes = [u64, i64, i64]
    lib.bfCudaSyntheticDiagValue.restype = C.c_double
    lib.bfCudaPlanGetStageEntries.argtypes = [vp, i32, C.POINTER(i64), i64, C.POINTER(i64)]
    lib.bfCudaPlanGetStageEntries.restype = i32
    lib.bfCudaDebugCompile.argtypes = [C.POINTER(PlanDesc), i32, i32, C.POINTER(C.POINTER(DebugTables))]
    lib.bfCudaDebugCompile.restype = i32
    lib.bfCudaDebugFree.argtypes = [C.POINTER(DebugTables)]
    lib.bfCudaDebugFree.restype = None
    lib.bfCudaPlanSetPerms.argtypes = [vp, C.POINTER(i64), C.POINTER(i64)]
    lib.bfCudaPlanSetPerms.restype = i32
    lib.bfCudaApplyDevice.argtypes = [vp, i32, vp, i64, i64, vp, i64, vp]
    lib.bfCudaApplyDevice.restype = i32
    lib.bfCudaApplyHost.argtypes = [vp, i32, vp, i64, i64, vp, i64]
    lib.bfCudaApplyHost.restype = i32
    lib.bfCudaPlanLastLaunchCount.argtypes = [vp]
    lib.bfCudaPlanLastLaunchCount.restype = i64
    lib.bfCudaPlanSetProfiling.argtypes = [vp, i32]
    lib.bfCudaPlanSetProfiling.restype = i32
    lib.bfCudaPlanGetStageTimes.argtypes = [vp, C.POINTER(C.c_float), i64, C.POINTER(i64)]
    lib.bfCudaPlanGetStageTimes.restype = i32
    lib.bfCudaPlanSave.argtypes = [vp, C.c_char_p]
    lib.bfCudaPlanSave.restype = i32
    lib.bfCudaPlanLoad.argtypes = [C.c_char_p, i32, C.POINTER(vp)]
    lib.bfCudaPlanLoad.restype = i32
    lib.bfCudaShardRankLoad.argtypes = [C.c_char_p, i32, C.POINTER(vp)]
    lib.bfCudaShardRankLoad.restype = i32
    lib.bfCudaRawOpsSave.argtypes = [C.POINTER(PlanDesc), C.c_char_p]
    lib.bfCudaRawOpsSave.restype = i32
    lib.bfCudaRawOpsLoad.argtypes = [C.c_char_p, C.POINTER(C.POINTER(PlanDesc))]
    lib.bfCudaRawOpsLoad.restype = i32
    lib.bfCudaRawOpsFree.argtypes = [C.POINTER(PlanDesc)]
    lib.bfCudaRawOpsFree.restype = None
    lib.bfCudaStatusString.argtypes = [i32]
    lib.bfCudaStatusString.restype = C.c_char_p
    lib.bfCudaLastErrorMessage.argtypes = []
    lib.bfCudaLastErrorMessage.restype = C.c_char_p
    lib.bfCudaMeasureFp64Peak.argtypes = [i32, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.bfCudaMeasureFp64Peak.restype = i32
    lib.bfCudaMeasureFp64PeakSustained.argtypes = [i32, C.c_double, C.POINTER(C.c_double)]
    lib.bfCudaMeasureFp64PeakSustained.restype = i32
    lib.bfCudaApplyStages.argtypes = [vp, i32, vp, i64, i64, vp, i64, i64, i64, vp]
    lib.bfCudaApplyStages.restype = i32
    lib.bfCudaPlanGetWorkspace.argtypes = [vp, i32, i64, C.POINTER(vp), C.POINTER(i64)]
    lib.bfCudaPlanGetWorkspace.restype = i32
    lib.bfCudaPlanGetWorkspaceLayout.argtypes = [vp, i32, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]
    lib.bfCudaPlanGetWorkspaceLayout.restype = i32
    lib.bfCudaPlanSetFused.argtypes = [vp, i32]
    lib.bfCudaPlanSetFused.restype = i32
    lib.bfCudaPlanGetFused.argtypes = [vp, i32]
    lib.bfCudaPlanGetFused.restype = i32
    lib.bfCudaPlanGetStream.argtypes = [vp]
    lib.bfCudaPlanGetStream.restype = vp
    lib.bfCudaSolveGMRES.argtypes = [vp, vp, vp, C.c_double, i64, vp, C.POINTER(i64), C.POINTER(C.c_double), C.POINTER(i64)]
    lib.bfCudaSolveGMRES.restype = i32
    lib.bfCudaGmresScratchBytes.argtypes = [i64]
    lib.bfCudaGmresScratchBytes.restype = i64
    lib.bfCudaGmresProject.argtypes = [vp, i64, i64, vp, i64, vp, vp, vp]
    lib.bfCudaGmresProject.restype = i32
    lib.bfCudaGmresSubtract.argtypes = [vp, i64, i64, vp, vp, i64, vp, vp, vp]
    lib.bfCudaGmresSubtract.restype = i32
    lib.bfCudaGmresProjectStatic.argtypes = [vp, i64, i64, vp, i64, vp, vp]
    lib.bfCudaGmresProjectStatic.restype = i32
    lib.bfCudaGmresScale.argtypes = [vp, i64, vp, vp, vp, vp]
    lib.bfCudaGmresScale.restype = i32
    PD = C.POINTER(PlanDesc)
    lib.bfCudaShardDescsCreate.argtypes = [PD, i32, C.POINTER(C.c_double), C.POINTER(vp)]
    lib.bfCudaShardDescsCreate.restype = i32
    lib.bfCudaShardDescsGet.argtypes = [vp, i32]
    lib.bfCudaShardDescsGet.restype = PD
    lib.bfCudaShardDescsGetLayout.argtypes = [vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i32), C.POINTER(i64), C.POINTER(i64)]
    lib.bfCudaShardDescsGetLayout.restype = i32
    lib.bfCudaShardDescsFree.argtypes = [vp]
    lib.bfCudaShardDescsFree.restype = None
    lib.bfCudaShardRankCreate.argtypes = [PD, C.POINTER(vp)]
    lib.bfCudaShardRankCreate.restype = i32
    lib.bfCudaShardRankDestroy.argtypes = [vp]
    lib.bfCudaShardRankDestroy.restype = i32
    lib.bfCudaShardRankGetPlan.argtypes = [vp]
    lib.bfCudaShardRankGetPlan.restype = vp
    lib.bfCudaShardRankSentRows.argtypes = [vp]
    lib.bfCudaShardRankSentRows.restype = i64
    lib.bfCudaShardRankGetHandle.argtypes = [vp, C.POINTER(ShardHandle)]
    lib.bfCudaShardRankGetHandle.restype = i32
    lib.bfCudaShardRankConnect.argtypes = [vp, C.POINTER(ShardHandle)]
    lib.bfCudaShardRankConnect.restype = i32
    lib.bfCudaShardRankApplyDevice.argtypes = [vp, vp, vp, vp]
    lib.bfCudaShardRankApplyDevice.restype = i32
    lib.bfCudaShardRankApplyStages.argtypes = [vp, vp, vp, i64, i64, vp]
    lib.bfCudaShardRankApplyStages.restype = i32
    lib.bfCudaCommInit.argtypes = [i32, C.POINTER(i32), C.POINTER(vp)]
    lib.bfCudaCommInit.restype = i32
    lib.bfCudaCommDestroy.argtypes = [vp]
    lib.bfCudaCommDestroy.restype = i32
    lib.bfCudaCommSize.argtypes = [vp]
    lib.bfCudaCommSize.restype = i32
    lib.bfCudaPlanShard.argtypes = [PD, vp, i32, C.POINTER(C.c_double), C.POINTER(vp)]
    lib.bfCudaPlanShard.restype = i32
    lib.bfCudaShardedPlanDestroy.argtypes = [vp]
    lib.bfCudaShardedPlanDestroy.restype = i32
    lib.bfCudaShardedPlanGetInfo.argtypes = [vp, C.POINTER(ShardedInfo)]
    lib.bfCudaShardedPlanGetInfo.restype = i32
    lib.bfCudaShardedPlanGetRankPlan.argtypes = [vp, i32]
    lib.bfCudaShardedPlanGetRankPlan.restype = vp
    lib.bfCudaShardedApplyHost.argtypes = [vp, i32, vp, i64, i64, vp, i64]
    lib.bfCudaShardedApplyHost.restype = i32
    lib.bfCudaShardedApplyDevice.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    lib.bfCudaShardedApplyDevice.restype = i32
    lib.bfCudaShardedSynchronize.argtypes = [vp]
    lib.bfCudaShardedSynchronize.restype = i32
    lib.bfCudaShardedRankStream.argtypes = [vp, i32]
    lib.bfCudaShardedRankStream.restype = vp
    lib.bfCudaAbiVersion.argtypes = []
    lib.bfCudaAbiVersion.restype = i32
    if path is None:
        _lib = lib
    return lib


def check(status: int, where: str):
    if status != OK:
        lib = load_library()
        detail = lib.bfCudaLastErrorMessage()
        raise BfCudaError(status, where, (detail or b"").decode(errors="replace"))
