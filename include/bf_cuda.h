/* bf_cuda.h — C ABI of libbfcuda.so: the B200 (sm_100a) butterfly-apply engine.
 *
 * This is the drop-in boundary for ONE path of sampotter/butterfly: applying a
 * precomputed butterfly factorization (a BfMat tree of MatProduct / MatBlockDiag /
 * MatBlockCoo / MatBlockDense nodes with dense, identity and diagonal leaves) to a block of
 * right-hand sides.  Everything here is `extern "C"`, plain pointers and sizes; there are no
 * reference types and no torch types in any signature.  The reference-facing half (walking a
 * `BfMat *` and swapping its vtable) is in bf_cuda_attach.h and is built on top of this file.
 *
 * Reference interfaces replaced (paths relative to the reference tree):
 *   bfMatMul / bfMatMulVec / bfMatRmul / bfMatRmulVec      src/mat.c:183-201
 *   bfMatProductMul / MulVec / Rmul / RmulVec              src/mat_product.c:211-345
 *   bfMatBlockDiagMul / MulVec / RmulVec                   src/mat_block_diag.c:370-505
 *   bfMatBlockCooMul / MulVec / RmulVec                    src/mat_block_coo.c:382-520
 *   bfMatBlockDenseMul / MulVec / Rmul / RmulVec           src/mat_block_dense.c:512-758
 *   bfMatDenseComplexMul (cblas_zgemm)                     src/mat_dense_complex.c:1024-1051,1704-1765
 *   bfMatDenseRealMulVec / RmulVec (cblas_dgemv)           src/mat_dense_real.c:1340-1441
 *   bfMatIdentityMul / MulVec / RmulVec                    src/mat_identity.c:149-199
 *   bfMatDiagRealMulVec                                    src/mat_diag_real.c:168
 *   bfMatDenseComplexPermuteRows / bfVecRealPermute        src/mat_dense_complex.c:741-774, src/vec_real.c:312-330
 *
 * Model.  The recursive operator is flattened ONCE into a list of "raw ops".  Each raw op is
 * one leaf of the tree together with where its operand comes from and where its result is
 * accumulated:
 *
 *        out[outBuf][outOff : outOff + m, :]  +=  op(block) * in[inBuf][inOff : inOff + k, :]
 *
 * Buffers are "virtual vectors" (row blocks of an n x nrhs matrix): buffer 0 is the operand X,
 * buffer 1 is the result Y, buffers >= 2 are the intermediates a MatProduct creates between its
 * factors (src/mat_product.c:225-238).  Every buffer starts at zero (bfMatZerosLike,
 * src/mat_block_coo.c:401; rows of a MatBlockDiag result that no block covers are
 * uninitialised in the reference, src/mat_block_diag.c:384 — we define them as zero).
 * bfCudaPlanCreate compiles the raw ops into a device-resident, stage-major packed layout
 *  and owns all device memory.
 *
 * Errors: every function returns a BfCudaStatus; nothing aborts.  The reference's convention
 * (global error code + NULL, src/error.c:5-24) is reproduced by the attach layer.
 *
 * There is NO CPU fallback: if no CUDA device is usable, bfCudaPlanCreate fails with
 * BF_CUDA_ERROR_NO_DEVICE.
 */
#ifndef BF_CUDA_H
#define BF_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BF_CUDA_ABI_VERSION 2

typedef enum BfCudaStatus {
  BF_CUDA_OK = 0,
  BF_CUDA_ERROR_INVALID_ARGUMENT = 1, /* ~ BF_ERROR_INVALID_ARGUMENTS (include/bf/error.h) */
  BF_CUDA_ERROR_NOT_IMPLEMENTED = 2,  /* ~ BF_ERROR_NOT_IMPLEMENTED */
  BF_CUDA_ERROR_OUT_OF_MEMORY = 3,    /* ~ BF_ERROR_MEMORY_ERROR */
  BF_CUDA_ERROR_RUNTIME = 4,          /* ~ BF_ERROR_RUNTIME_ERROR (a CUDA call failed) */
  BF_CUDA_ERROR_NO_DEVICE = 5,
  BF_CUDA_ERROR_IO = 6,               /* ~ BF_ERROR_FILE_ERROR */
  BF_CUDA_ERROR_INCOMPATIBLE_SHAPES = 7 /* ~ BF_ERROR_INCOMPATIBLE_SHAPES */
} BfCudaStatus;

typedef enum BfCudaDtype {
  BF_CUDA_REAL_F64 = 0,   /* BfReal = double              (include/bf/def.h:23-34) */
  BF_CUDA_COMPLEX_F64 = 1 /* BfComplex = double _Complex, interleaved (re, im) */
} BfCudaDtype;

/* Which operator is applied.  N: Y = A X (bfMatMul / bfMatMulVec).  T: Y = A^T X
 * (bfMatRmulVec on the real path, src/mat_product.c:312-345).  H: Y = A^H X (what
 * bfMatTranspose means for complex dense leaves: it sets TRANS and CONJ,
 * src/mat_dense_complex.c:1475-1478). */
typedef enum BfCudaOp { BF_CUDA_OP_N = 0, BF_CUDA_OP_T = 1, BF_CUDA_OP_H = 2 } BfCudaOp;

typedef enum BfCudaLeafKind {
  BF_CUDA_LEAF_DENSE = 0,    /* BfMatDenseComplex / BfMatDenseReal: m x k values, row stride ld */
  BF_CUDA_LEAF_IDENTITY = 1, /* BfMatIdentity: m == k, no values (src/mat_identity.c:149-199) */
  BF_CUDA_LEAF_DIAG = 2      /* BfMatDiagReal: min(m,k) REAL values d[i] (src/mat_diag_real.c:168) */
} BfCudaLeafKind;

/* Leaf property bits, same meaning as BfMatProps TRANS / CONJ (include/bf/mat.h:30-39).  When
 * TRANS is set, `val` holds the k x m STORED matrix and the leaf acts as its transpose.  As in
 * the reference's getCblasTranspose (src/mat_dense_complex.c:27-35) a complex leaf with either
 * bit set acts as the CONJUGATE transpose. */
#define BF_CUDA_LEAF_TRANS 2u
#define BF_CUDA_LEAF_CONJ 4u
/* Block-row sharding: the leaf belongs to the operator's structure — stages,
 * workspace layout and row addresses are computed from ALL leaves, identically on every rank — but
 * another rank applies it: no items, tiles or pool bytes are generated for it here, and the rows it
 * writes are expected to arrive through the caller's exchange (bfCudaApplyStages). */
#define BF_CUDA_LEAF_REMOTE 8u
/* Sharded descriptors: bits 8..15 of `flags` hold the rank that applies the leaf (of a REMOTE leaf: which other
 * rank; of a local one: this rank).  The plan compiler derives from them which rows a stage kernel also stores
 * into which peer's workspace (BfCudaPlanDesc.shardWorld > 1). */
#define BF_CUDA_LEAF_OWNER_SHIFT 8
#define BF_CUDA_LEAF_OWNER(flags) (((flags) >> BF_CUDA_LEAF_OWNER_SHIFT) & 0xFFu)

enum { BF_CUDA_BUF_X = 0, BF_CUDA_BUF_Y = 1, BF_CUDA_BUF_FIRST_TEMP = 2 };

/* One leaf of the flattened operator.  m x k is the shape the leaf ACTS with. */
typedef struct BfCudaRawOp {
  int32_t kind;   /* BfCudaLeafKind */
  uint32_t flags; /* BF_CUDA_LEAF_TRANS | BF_CUDA_LEAF_CONJ */
  int64_t inBuf;  /* virtual buffer read */
  int64_t inOff;  /* first row read */
  int64_t outBuf; /* virtual buffer accumulated into */
  int64_t outOff; /* first row written */
  int64_t m;      /* rows produced */
  int64_t k;      /* rows consumed */
  int64_t ld;     /* row stride of the STORED matrix, in elements (>= stored column count) */
  const void *val; /* host pointer to stored element (0,0); NULL = "no host values": the pool is
                      left for bfCudaPlanFillSynthetic (benchmarks on synthetic factors) */
} BfCudaRawOp;

#define BF_CUDA_BUILD_KEEP_LEAVES 4

typedef struct BfCudaPlanDesc {
  int32_t abiVersion;     /* BF_CUDA_ABI_VERSION */
  int32_t dtype;          /* BfCudaDtype of every dense leaf and of X, Y */
  int64_t numRows;        /* rows of A  == rows of buffer 1 (Y) */
  int64_t numCols;        /* cols of A  == rows of buffer 0 (X) */
  int64_t numBufs;        /* >= 2 */
  const int64_t *bufRows; /* [numBufs] rows of every virtual buffer */
  int64_t numOps;
  const BfCudaRawOp *ops; /* [numOps], in the reference's traversal order: this order IS the
                             summation order used for every output row (deterministic) */
  int32_t device;         /* CUDA device ordinal, -1 = current device */
  int32_t buildTranspose; /* bit 0: also compile A^T (OP_T); bit 1: also compile A^H (OP_H,
                             complex only).  Each extra direction holds its own packed pool.
                             bit 2 (BF_CUDA_BUILD_KEEP_LEAVES): compile the leaf list exactly as given — no
                             merging of stacked small leaves, no elision of identity copies — for callers
                             that reason about the plan's stages and buffers leaf by leaf */
  int32_t shardWorld;     /* 0 / 1: the whole operator on one device.  > 1: one rank's share of a block-row
                             sharded operator (bfCudaShardDesc builds these): leaves of other ranks carry
                             BF_CUDA_LEAF_REMOTE, every leaf its owner (BF_CUDA_LEAF_OWNER); forward only */
  int32_t shardRank;      /* this rank, 0 <= shardRank < shardWorld */
} BfCudaPlanDesc;

typedef struct BfCudaPlan BfCudaPlan; /* opaque */

typedef struct BfCudaPlanInfo {
  int32_t dtype;
  int32_t device;
  int64_t numRows, numCols;
  int64_t numOps;            /* raw ops */
  int64_t numEntries;        /* E = sum over dense leaves of m*k */
  int64_t numStagesN;        /* kernel stages per forward apply */
  int64_t numStagesT;        /* ... per transposed apply (0 if not built) */
  int64_t numItemsN;         /* output segments over all forward stages */
  int64_t numTilesN;         /* block pieces over all forward stages */
  int64_t poolBytes;         /* packed factor values resident in HBM */
  int64_t tableBytes;        /* index tables resident in HBM */
  int64_t workspaceRowsN;    /* rows of intermediate workspace per RHS column (forward) */
  int64_t workspaceRowsT;
  int64_t maxTileK;          /* widest tile after splitting */
  int64_t maxItemRows;       /* tallest output segment after splitting */
} BfCudaPlanInfo;

/* ---- life cycle -------------------------------------------------------------------------- */

/* Compile + upload.  Host value pointers in `desc` are only read during this call. */
BfCudaStatus bfCudaPlanCreate(const BfCudaPlanDesc *desc, BfCudaPlan **plan);
BfCudaStatus bfCudaPlanDestroy(BfCudaPlan *plan);
BfCudaStatus bfCudaPlanGetInfo(const BfCudaPlan *plan, BfCudaPlanInfo *info);

/* Fill the value pool on the device with a reproducible pseudo-random pattern; used by
 * benchmarks whose raw ops carry val == NULL (synthetic factors of the named size: real structure,
 * synthetic values).  The value of STORED element (row, col) of dense raw op `opIndex` is
 * bfCudaSyntheticValue(seed, opIndex, row * storedCols + col, part, k): uniform in
 * [-2^-round(log2(k)/2), +...) with k the op's contraction length, so that repeated levels stay
 * O(1).  Diagonal leaves get bfCudaSyntheticDiagValue(seed, opIndex, i) in [0.5, 1.5).  Both are
 * integer hashes followed by one exact multiply: a CPU checker regenerates identical bits. */
BfCudaStatus bfCudaPlanFillSynthetic(BfCudaPlan *plan, uint64_t seed);
double bfCudaSyntheticValue(uint64_t seed, int64_t opIndex, int64_t storedLinear, int32_t part,
                            int64_t k);
double bfCudaSyntheticDiagValue(uint64_t seed, int64_t opIndex, int64_t i);

/* Optional caller-side permutations fused into the first gather / last scatter
 * (reference callers do them as separate passes: examples/simple/helm2_bie.c:70-81,139).
 * Semantics follow bfVecRealPermute / bfMatDenseComplexPermuteRows (out[perm[i]] = in[i]):
 *   inPerm  (length numCols, may be NULL): the operand the plan sees is P_in applied to X,
 *            i.e. Xp[inPerm[i], :] = X[i, :]
 *   outPerm (length numRows, may be NULL): the caller receives Yp with Yp[outPerm[i], :] = Y[i, :]
 * Both must be bijections; checked, BF_CUDA_ERROR_INVALID_ARGUMENT otherwise. */
BfCudaStatus bfCudaPlanSetPerms(BfCudaPlan *plan, const int64_t *inPerm, const int64_t *outPerm);

/* ---- apply --------------------------------------------------------------------------------- */

/* Device-resident apply (the timed entry point).  d_x: (numCols x nrhs) for OP_N, (numRows x
 * nrhs) for OP_T / OP_H; row-major, row strides ldx / ldy in ELEMENTS (>= nrhs).  `stream` is a
 * cudaStream_t passed as void* (NULL = the plan's own stream).  Asynchronous w.r.t. the host. */
BfCudaStatus bfCudaApplyDevice(BfCudaPlan *plan, int32_t op, const void *d_x, int64_t ldx,
                               int64_t nrhs, void *d_y, int64_t ldy, void *stream);

/* One RHS column, whole operator: bfCudaApplyDevice runs the apply as ONE cooperative kernel launch over
 * "fused" tables — a queue-major copy of the stage tables with a flat workspace, in which the grid-wide
 * barrier between butterfly levels is replaced by per-buffer readiness counters.  Results are
 * bit-identical to the per-stage launches (same items, same summation order).  By default the single launch is
 * used from a pool size on where it measures faster (24 GiB: 0.942 -> 0.961 of the HBM peak at N = 524 288; on
 * small operators a stage is a few tiles per warp and the PDL-chained per-stage launches win).
 * bfCudaPlanSetFused(plan, 0 / 1 / -1) = never / always / by size (A/B measurements; the environment variable
 * BF_CUDA_FUSED=0/1 does the same for every plan); profiled applies (bfCudaPlanSetProfiling) and
 * bfCudaApplyStages always use per-stage launches.  bfCudaPlanGetFused: 1 if the next 1-RHS apply will be a
 * single launch. */
BfCudaStatus bfCudaPlanSetFused(BfCudaPlan *plan, int32_t enable);
int32_t bfCudaPlanGetFused(const BfCudaPlan *plan, int32_t op);

/* Stage-by-stage apply for callers that interleave an exchange between stages (block-row sharding:
 * every rank runs its own leaves of stage s, then rows of the workspace produced for other ranks
 * are exchanged before stage s + 1).  Runs stages [stageBegin, stageEnd) of direction `op`; the
 * operand is staged when stageBegin == 0.  nrhs must fit one workspace chunk (NOT_IMPLEMENTED
 * otherwise).  bfCudaPlanGetWorkspace gives the device workspace those stages read and write: row r
 * of the workspace starts at ws + r * ld * elemSize; valid until the next apply with a larger nrhs.
 * bfCudaPlanGetWorkspaceLayout: rows [0, arena0) and [arena0, arena0 + arena1) are the two
 * ping-pong arenas (a row written at stage t is read at stage t + 1 only), the remaining
 * `dedicated` rows belong to buffers written over several stages and are never reused. */
BfCudaStatus bfCudaApplyStages(BfCudaPlan *plan, int32_t op, const void *d_x, int64_t ldx, int64_t nrhs,
                               void *d_y, int64_t ldy, int64_t stageBegin, int64_t stageEnd, void *stream);
BfCudaStatus bfCudaPlanGetWorkspace(BfCudaPlan *plan, int32_t op, int64_t nrhs, void **d_ws, int64_t *ld);
BfCudaStatus bfCudaPlanGetWorkspaceLayout(const BfCudaPlan *plan, int32_t op, int64_t *arena0Rows,
                                          int64_t *arena1Rows, int64_t *dedicatedRows);

/* ---- one operator over several GPUs -------------------------------------------------------
 * The reference is single-process, single-thread C; its callers loop over bfMatMul (src/linalg.c:125,155;
 * examples/multiple_scattering/multiple_scattering_context.c:776-794).  Nothing in it shards an operator, so
 * these entry points replace no reference interface: they put several GPUs behind the SAME calls.
 *
 * Block rows.  bfCudaShardDescsCreate decides which rank applies which leaf (contiguous row ranges of Y and of
 * every intermediate level; leaves taller than 2048 rows and diagonal terms are cut at the boundaries) and
 * returns one descriptor per rank: the same leaves everywhere, those of other ranks flagged
 * BF_CUDA_LEAF_REMOTE.  bfCudaShardDescsGetLayout: yCuts[world + 1] (rank r owns rows [yCuts[r], yCuts[r+1]) of
 * the result), and per leaf of the sharded list its owner and where it comes from in the original list
 * (originOp, originRow: per-leaf side data such as Helmholtz recipes follows the cuts through these).  Any of
 * the output pointers may be NULL. */
typedef struct BfCudaShardDescs BfCudaShardDescs;
BfCudaStatus bfCudaShardDescsCreate(const BfCudaPlanDesc *desc, int32_t world, const double *fractions,
                                    BfCudaShardDescs **out);
const BfCudaPlanDesc *bfCudaShardDescsGet(const BfCudaShardDescs *descs, int32_t rank);
BfCudaStatus bfCudaShardDescsGetLayout(const BfCudaShardDescs *descs, int64_t *numOps, int64_t *yCuts,
                                       int32_t *owner, int64_t *originOp, int64_t *originRow);
void bfCudaShardDescsFree(BfCudaShardDescs *descs);

/* One rank at run time: its plan (bfCudaShardRankGetPlan: fill its values with bfCudaPlanBuildHelm2 /
 * bfCudaPlanFillSynthetic, or let the descriptor carry host values), a workspace and a flag array its peers can
 * store into.  The stage kernels write the rows another rank reads straight into that rank's workspace (peer
 * memory over NVLink) and post one flag per (stage, peer); the reader's next stage spins on the flag.  There is no
 * collective and no host round trip between stages.  Set-up: every rank exports a BfCudaShardHandle, the caller
 * moves the handles between ranks (memcpy in one process; an all_gather of sizeof(BfCudaShardHandle) bytes
 * between processes) and every rank connects to all of them: same process = the raw pointers (peer access is
 * enabled), another process = the CUDA IPC handles.  One column per apply. */
typedef struct BfCudaShardRank BfCudaShardRank;
typedef struct BfCudaShardHandle {
  int64_t pid;            /* process that owns the memory */
  int32_t device, rank;
  int64_t wsBytes;
  void *wsPtr, *flagsPtr; /* valid in the owning process only */
  int32_t hasIpc, reserved;
  unsigned char wsIpc[64], flagsIpc[64]; /* cudaIpcMemHandle_t */
} BfCudaShardHandle;
BfCudaStatus bfCudaShardRankCreate(const BfCudaPlanDesc *rankDesc, BfCudaShardRank **out);
BfCudaStatus bfCudaShardRankLoad(const char *path, int32_t device, BfCudaShardRank **out); /* a rank plan saved with bfCudaPlanSave */
BfCudaStatus bfCudaShardRankDestroy(BfCudaShardRank *rank);
BfCudaPlan *bfCudaShardRankGetPlan(BfCudaShardRank *rank);
int64_t bfCudaShardRankSentRows(const BfCudaShardRank *rank); /* workspace rows pushed to peers per apply */
BfCudaStatus bfCudaShardRankGetHandle(BfCudaShardRank *rank, BfCudaShardHandle *handle);
BfCudaStatus bfCudaShardRankConnect(BfCudaShardRank *rank, const BfCudaShardHandle *allRanks);
/* d_x: the FULL operand (numCols elements, dense, 16-byte aligned) on this rank's device; d_y: numRows
 * elements, of which this rank writes the rows it owns.  Asynchronous.  ...Stages runs stages [stageBegin,
 * stageEnd) (stageEnd < 0: to the end); stageBegin == 0 starts a new apply.  A driver of several ranks in one
 * thread launches stage-major so that no device starts late. */
BfCudaStatus bfCudaShardRankApplyDevice(BfCudaShardRank *rank, const void *d_x, void *d_y, void *stream);
BfCudaStatus bfCudaShardRankApplyStages(BfCudaShardRank *rank, const void *d_x, void *d_y, int64_t stageBegin,
                                        int64_t stageEnd, void *stream);

/* All ranks in ONE process, driven by one host thread: what a reference program uses (bfCudaMatAttachSharded in
 * bf_cuda_attach.h puts it behind bfMatMul).  bfCudaCommInit: the devices to use (NULL: 0 .. nGpus - 1; an
 * ordinal may repeat — ranks sharing a GPU run one after the other, which is how the single-GPU test exercises
 * the protocol); one stream per rank.  bfCudaPlanShard: BF_CUDA_SHARD_BLOCK_ROWS as above (forward only, one
 * column at a time: a block of columns is applied column by column); BF_CUDA_SHARD_RHS_COLS replicates the plan
 * (desc->buildTranspose honoured) and gives every device a slice of the columns — the mode for wide blocks. */
typedef struct BfCudaComm BfCudaComm;
typedef struct BfCudaShardedPlan BfCudaShardedPlan;
typedef enum BfCudaShardMode { BF_CUDA_SHARD_RHS_COLS = 0, BF_CUDA_SHARD_BLOCK_ROWS = 1 } BfCudaShardMode;
typedef struct BfCudaShardedInfo {
  int32_t world, mode;
  int64_t numRows, numCols;
  int64_t numEntries;           /* E of the whole operator */
  int64_t maxRankEntries;       /* largest per-rank share (block rows) */
  int64_t poolBytesMax;         /* largest per-device factor pool */
  int64_t sentRowsPerApply;     /* workspace rows pushed between GPUs per apply, all ranks (block rows) */
  int64_t numStages;
  int64_t yCuts[17];            /* block rows: row ranges of the result */
} BfCudaShardedInfo;
BfCudaStatus bfCudaCommInit(int32_t nGpus, const int32_t *devices, BfCudaComm **comm);
BfCudaStatus bfCudaCommDestroy(BfCudaComm *comm);
int32_t bfCudaCommSize(const BfCudaComm *comm);
BfCudaStatus bfCudaPlanShard(const BfCudaPlanDesc *desc, BfCudaComm *comm, int32_t mode, const double *fractions,
                             BfCudaShardedPlan **plan);
BfCudaStatus bfCudaShardedPlanDestroy(BfCudaShardedPlan *plan);
BfCudaStatus bfCudaShardedPlanGetInfo(const BfCudaShardedPlan *plan, BfCudaShardedInfo *info);
BfCudaPlan *bfCudaShardedPlanGetRankPlan(BfCudaShardedPlan *plan, int32_t rank);
/* Host buffers, synchronous: X to every device that needs it, apply, the result rows / columns back. */
BfCudaStatus bfCudaShardedApplyHost(BfCudaShardedPlan *plan, int32_t op, const void *x, int64_t ldx, int64_t nrhs,
                                    void *y, int64_t ldy);
/* Block rows, device resident (the timed entry point): d_x[r] = the full operand on rank r's device, d_y[r] =
 * numRows elements there (rank r writes its own rows).  Asynchronous on the ranks' streams;
 * bfCudaShardedSynchronize waits for all of them.  bfCudaShardedRankStream: rank r's stream (cudaStream_t). */
BfCudaStatus bfCudaShardedApplyDevice(BfCudaShardedPlan *plan, const void *const *d_x, void *const *d_y);
BfCudaStatus bfCudaShardedSynchronize(BfCudaShardedPlan *plan);
void *bfCudaShardedRankStream(BfCudaShardedPlan *plan, int32_t rank);

/* Host-buffer apply: H2D of X, apply, D2H of Y, synchronous.  x / y are ordinary host memory
 * (the reference's malloc'ed BfMatDenseComplex::data / BfVecReal::data). */
BfCudaStatus bfCudaApplyHost(BfCudaPlan *plan, int32_t op, const void *x, int64_t ldx,
                             int64_t nrhs, void *y, int64_t ldy);

/* The plan's own CUDA stream (a cudaStream_t as void*): what a NULL `stream` argument means. */
void *bfCudaPlanGetStream(const BfCudaPlan *plan);

/* Device-resident GMRES around the apply: solves A x = b for ONE complex right-hand side with the
 * algorithm of bfSolveGMRES (src/linalg.c:47-318) — modified Gram-Schmidt Arnoldi, the reference's
 * Givens rotations (src/vec_complex.c:155-168,230-290), convergence when |s[j+1]| / ||r0|| < tol,
 * and, like the reference, the iterate rebuilt from the first *numIter basis vectors (the column on
 * which the test fires is not counted, src/linalg.c:224-238).  No preconditioner.  b, x0 (may be
 * NULL = zero) and x are host arrays of numRows complex doubles; the Krylov basis stays in HBM.
 * relResidual (may be NULL) receives the last |s[j+1]| / ||r0||, numApplies the operator applies. */
BfCudaStatus bfCudaSolveGMRES(BfCudaPlan *plan, const void *b, const void *x0, double tol,
                              int64_t maxNumIter, void *x, int64_t *numIter, double *relResidual,
                              int64_t *numApplies);

/* The same solve with a LEFT preconditioner (bfSolveGMRES's `M` argument, src/linalg.c:42-49,131,159):
 * `precond(ctx, in, out, n)` must write out = M^{-1} in for host arrays of n complex doubles and return
 * 0.  It is called once for the initial residual and once per iteration, between the apply and the
 * Gram-Schmidt pass; the vector crosses PCIe once each way per call (the reference's preconditioners
 * are host LU solves).  relResidual is then the preconditioned residual, as in the reference. */
typedef int (*BfCudaPrecondFn)(void *ctx, const void *in, void *out, int64_t n);
BfCudaStatus bfCudaSolveGMRESPrecond(BfCudaPlan *plan, const void *b, const void *x0, double tol,
                                     int64_t maxNumIter, void *x, int64_t *numIter,
                                     double *relResidual, int64_t *numApplies, BfCudaPrecondFn precond,
                                     void *precondCtx);

/* Local halves of a Gram-Schmidt pass for a solve whose Krylov basis is distributed by rows over
 * several GPUs (butterfly_b200/gmres.py; the caller all_reduces `out` / `nrm2Out` between them).
 * All pointers are DEVICE pointers to complex doubles, `stream` a cudaStream_t, `scratch` at least
 * bfCudaGmresScratchBytes(m) bytes.  Sums are two-stage in a fixed order (bitwise reproducible).
 *   Project : out[i] = sum_k conj(V[i*ldv + k]) w[k], i < m (zdotc as src/mat_dense_complex.c:115-121);
 *             out[m] = (sum_k |w[k]|^2, 0)
 *   Subtract: w -= sum_{i<m} h[i] V[i];  nrm2Out[0] = (sum_k |w_new[k]|^2, 0)   (src/linalg.c:172-190)
 *   Scale   : dst = w / sqrt(nrm2[0].re), also written to dst2 when non-NULL      (src/linalg.c:196-202) */
int64_t bfCudaGmresScratchBytes(int64_t m);
BfCudaStatus bfCudaGmresProject(const void *V, int64_t ldv, int64_t m, const void *w, int64_t n,
                                void *out, void *scratch, void *stream);
BfCudaStatus bfCudaGmresSubtract(const void *V, int64_t ldv, int64_t m, const void *h, void *w,
                                 int64_t n, void *nrm2Out, void *scratch, void *stream);
BfCudaStatus bfCudaGmresScale(const void *w, int64_t n, const void *nrm2, void *dst, void *dst2,
                              void *stream);

/* Allocation-free variant of bfCudaGmresProject.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments and results as bfCudaGmresProject, minus `scratch`: the
 * partial sums live in a static device array inside the library (606 208 bytes, held by every GPU that loads
 * the library until the process exits), looked up once per device with cudaGetSymbolAddress; the call issues no
 * cudaMalloc / cudaFree / allocating library routine.  Limits and rules:
 *   - m <= BF_CUDA_GMRES_PROJECT_STATIC_MAX_M basis vectors; a larger m is refused with
 *     BF_CUDA_ERROR_INVALID_ARGUMENT and a message in bfCudaLastErrorMessage (use bfCudaGmresProject then);
 *   - V, w, out must live on the CURRENT device;
 *   - every call on a device shares the one array: calls that overlap on the GPU (different streams, no
 *     dependency between them) RACE — serialise them.  The array keeps the last call's contents; nothing has to
 *     be cleared, every slot read is written earlier in the same call. */
#define BF_CUDA_GMRES_PROJECT_STATIC_MAX_M 62
BfCudaStatus bfCudaGmresProjectStatic(const void *V, int64_t ldv, int64_t m, const void *w, int64_t n,
                                      void *out, void *stream);

/* Number of kernel launches the last apply on this plan issued (bench.py's gpu_launches). */
int64_t bfCudaPlanLastLaunchCount(const BfCudaPlan *plan);

/* Per-stage device time of the last bfCudaApplyDevice call made with profiling enabled.
 * ms must hold numStages floats. */
BfCudaStatus bfCudaPlanSetProfiling(BfCudaPlan *plan, int32_t enable);
BfCudaStatus bfCudaPlanGetStageTimes(BfCudaPlan *plan, float *ms, int64_t capacity,
                                     int64_t *numStages);
/* Dense entries (sum of m*k) applied by every stage of direction `op`: the per-stage share of E. */
BfCudaStatus bfCudaPlanGetStageEntries(const BfCudaPlan *plan, int32_t op, int64_t *entries,
                                       int64_t capacity, int64_t *numStages);

/* ---- persistence ----------------------------------------------------------------------------
 * A flat little-endian file holding the raw ops + values, so a factorization that took the host
 * an hour to build is built once (the reference has bfMatDump writers but no reader,
 * src/mat.c:67-73). */
BfCudaStatus bfCudaRawOpsSave(const BfCudaPlanDesc *desc, const char *path);
/* Loads into one malloc'ed arena; free with bfCudaRawOpsFree. */
BfCudaStatus bfCudaRawOpsLoad(const char *path, BfCudaPlanDesc **desc);
void bfCudaRawOpsFree(BfCudaPlanDesc *desc);

/* The COMPILED plan — stage tables, single-launch tables, fused permutations and the packed factor pool as it sits
 * in HBM, every built direction — in one file.  Covers what bfCudaRawOpsSave cannot: operators whose values only
 * ever existed on the GPU (bfCudaPlanBuildHelm2) and the per-rank plans of a block-row sharded operator
 * (bfCudaPlanSave(bfCudaShardRankGetPlan(rank), path) / bfCudaShardRankLoad).  Reference anchors: the Dump
 * writers that have no reader, src/mat.c:67-73, src/mat_product.c:123-140, src/mat_block_coo.c:260-289.  The pool
 * moves through two pinned 64 MiB buffers (disk and PCIe overlap).  A loaded plan applies bit-identically to the
 * saved one on any sm_100 device (`device` = -1: the current one); it holds no per-leaf data, so
 * bfCudaPlanFillSynthetic / bfCudaPlanBuildHelm2 on it return NOT_IMPLEMENTED.  Truncated or corrupted files are
 * refused (BF_CUDA_ERROR_IO): every index the kernels would follow is checked against the file's own sizes. */
BfCudaStatus bfCudaPlanSave(BfCudaPlan *plan, const char *path);
BfCudaStatus bfCudaPlanLoad(const char *path, int32_t device, BfCudaPlan **plan);

/* ---- on-device construction of Helmholtz factor values ----------------------
 * The reference computes every leaf of a fac_helm2 operator with one of two builders:
 *   bfHelm2GetKernelMatrix       (src/helm2.c:282-319)  Z[i][j] = G(tgt_i, src_j)
 *   bfHelm2GetReexpansionMatrix  (src/helm2.c:321-365)  Z_equiv \ Z_orig, a truncated-SVD least
 *                                                       squares (src/mat_dense_complex.c:1768-1850)
 * A "recipe" is the argument list of one such call.  bfCudaPlanBuildHelm2 evaluates the recipes on
 * the GPU and writes the values straight into the plan's packed pool (every built direction), so the
 * O(N^1.4) value computation and the host->device upload of the factors both disappear.  The plan must
 * have been created from raw ops with val == NULL for the leaves named in opRecipe.
 *
 * Points and unit normals live in ONE pool of (x, y) pairs; recipes index it in pairs. */
typedef enum BfCudaHelm2Kind { BF_CUDA_HELM2_KERNEL = 0, BF_CUDA_HELM2_REEXPANSION = 1 } BfCudaHelm2Kind;

/* Values of the reference's BfLayerPotential (include/bf/layer_pot.h:29-43). */
enum { BF_CUDA_LAYER_POT_SINGLE = 1, BF_CUDA_LAYER_POT_PV_DOUBLE = 2,
       BF_CUDA_LAYER_POT_PV_NORMAL_DERIV_SINGLE = 3, BF_CUDA_LAYER_POT_COMBINED_FIELD = 5 };

typedef struct BfCudaHelm2Recipe {
  int32_t kind;       /* BfCudaHelm2Kind */
  int32_t layerPot;   /* BF_CUDA_LAYER_POT_* (BfHelm2::layerPot, include/bf/helm2.h:10-15) */
  double k;           /* wavenumber */
  double alpha[2], beta[2]; /* combined field: alpha S + beta D (re, im) */
  int64_t rows, cols; /* shape of the result: KERNEL numTgt x numSrc, REEXPANSION numEquiv x numSrc */
  int64_t src, numSrc;     /* (original) source points: first pair, count */
  int64_t tgt, numTgt;     /* target points */
  int64_t equiv, numEquiv; /* REEXPANSION: equivalent sources */
  int64_t srcNormals, tgtNormals, equivNormals; /* first pair of the unit normals, -1 = none */
} BfCudaHelm2Recipe;

typedef struct BfCudaHelm2Build {
  int64_t numPairs;
  const double *pairs;               /* [2 * numPairs] */
  int64_t numRecipes;
  const BfCudaHelm2Recipe *recipes;
  int64_t numOps;                    /* == raw ops of the plan */
  const int64_t *opRecipe;           /* [numOps] recipe whose result raw op i stores (whole matrix,
                                        ld == cols), or -1: leave the leaf's pool values alone */
  double truncationScale;            /* 0 or 1: the reference's rule, keep sigma_i >= max(m, n) eps sigma_0 + eps
                                        (src/mat_dense_complex.c:1796-1806); s > 0: s times that threshold.
                                        The operator's accuracy is sensitive to it: 4 gives a
                                        smaller error against the dense kernel than the reference's own build. */
  const int64_t *opRowOffset;        /* NULL, or [numOps]: leaf i holds the rows [opRowOffset[i], + its stored rows) of
                                        its recipe's result and all columns (leaves cut into row slabs for block-row
                                        sharding); NULL = every leaf is the whole matrix */
} BfCudaHelm2Build;

typedef struct BfCudaHelm2Stats {
  int64_t kernelRecipes, kernelEntries;   /* leaves evaluated directly into the pool */
  int64_t reexpRecipes, reexpEntries;     /* least-squares leaves */
  int64_t reexpSmall, reexpLarge;         /* solved in shared memory / in global scratch */
  int64_t maxSweeps;                      /* most Jacobi sweeps any problem needed */
  int64_t notConverged;                   /* problems that hit the sweep limit */
  double kernelMs, reexpMs, packMs;       /* device time: (unused), bfHelm2Pinv launches, bfHelm2Fill launch */
  double pinvClassMs[6];                  /* bfHelm2Pinv per launch class: Z_equiv and G both in shared memory
                                             (<= 32 / 72 / 110 / 225 KiB), Z_equiv only, neither */
  int64_t pinvClassCount[6];              /* distinct pseudo-inverses per class */
} BfCudaHelm2Stats;

BfCudaStatus bfCudaPlanBuildHelm2(BfCudaPlan *plan, const BfCudaHelm2Build *build, BfCudaHelm2Stats *stats);

/* The two builders on their own (tests, and callers that want the matrix rather than a plan): the result
 * of recipe `index`, row-major rows x cols complex doubles, to host memory. */
BfCudaStatus bfCudaHelm2Evaluate(const BfCudaHelm2Build *build, int64_t index, void *out, int32_t device);

/* ---- misc ---------------------------------------------------------------------------------- */
/* Measured fp64 ceilings of `device` (-1 = current) in TFLOP/s: a register-resident DFMA loop and a
 * DMMA (mma.sync.m8n8k4.f64) loop, ~20 ms each.  The roofline denominators of the many-RHS apply
 * (bench.py); not used by the apply itself. */
BfCudaStatus bfCudaMeasureFp64Peak(int32_t device, double *dfmaTflops, double *dmmaTflops);
/* The DMMA loop launched back to back for `seconds` (default 2), averaged over the second half: the
 * SUSTAINED fp64 tensor ceiling under the power cap — the denominator for a kernel timed inside a long
 * step (a 256-RHS apply runs for hundreds of milliseconds). */
BfCudaStatus bfCudaMeasureFp64PeakSustained(int32_t device, double seconds, double *dmmaTflops);
const char *bfCudaStatusString(BfCudaStatus status);
const char *bfCudaLastErrorMessage(void); /* thread-local, human-readable detail */
int32_t bfCudaAbiVersion(void);

#ifdef __cplusplus
}
#endif
#endif /* BF_CUDA_H */
