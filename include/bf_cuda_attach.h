/* bf_cuda_attach.h — the reference-facing half of the drop-in: walks a `BfMat *` tree of the
 * reference library and routes bfMatMul / bfMatMulVec / bfMatRmul / bfMatRmulVec on that object
 * to the GPU plan by swapping the object's public vtable pointer (include/bf/mat.h:187-188).
 *
 * This header needs the reference's headers on the include path (-I<reference>/include) and is
 * meant to be compiled INTO (or next to) libbutterfly; bf_cuda.h alone is reference-free.
 * INTEGRATION.md shows the two lines a maintainer adds.
 *
 * Interfaces replaced on an attached object (signatures identical, include/bf/mat.h:90-95):
 *   BfMat *Mul    (BfMat const *A, BfMat const *X)   src/mat_product.c:211, mat_block_dense.c:512 ...
 *   BfVec *MulVec (BfMat const *A, BfVec const *x)   src/mat_product.c:247, mat_block_dense.c:574 ...
 *   BfMat *Rmul   (BfMat const *A, BfMat const *X)   (computes X A)  src/mat_product.c:282
 *   BfVec *RmulVec(BfMat const *A, BfVec const *x)   (computes A^T x) src/mat_product.c:312
 *   BfSize NumBytes(BfMat const *A)                  unchanged value (forwards to the original)
 *   void   Delete (BfMat **A)                        destroys the plan, then the original Delete
 */
#ifndef BF_CUDA_ATTACH_H
#define BF_CUDA_ATTACH_H

#include <bf/mat.h>
#include <bf/perm.h>

#include "bf_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct BfCudaAttachOpts {
  int32_t device;         /* CUDA device ordinal, -1 = current */
  int32_t buildTranspose; /* also compile A^T / A^H (needed by Rmul / RmulVec) */
  uint32_t flattenFlags;  /* BF_CUDA_FLATTEN_* */
} BfCudaAttachOpts;

/* MatCooComplex leaves: by default several stored elements in one row behave as in the reference's
 * bfMatCooComplexMul, where each element OVERWRITES its result row (src/mat_coo_complex.c:248-251), so
 * only the last one counts.  With this flag they are summed, i.e. the sparse matrix is applied as the
 * matrix it stores. */
#define BF_CUDA_FLATTEN_COO_ACCUMULATE 1u

/* Flatten `A` into raw ops (bf_cuda.h).  Values are NOT copied: ops point into the leaves of
 * `A`, which must outlive the returned descriptor's use.  Supported nodes: MatProduct,
 * MatBlockDense, MatBlockDiag, MatBlockCoo, MatSum, MatDenseComplex, MatDenseReal, MatIdentity,
 * MatDiagReal, MatZero, and MatCooComplex (Kapur-Rokhlin correction terms inside a MatSum,
 * src/mat_block_dense.c:458-510), which becomes dense leaves in buffers the descriptor owns.  Anything
 * else (MatPerm has no Mul in the reference either, src/mat_perm.c:53-61) ->
 * BF_CUDA_ERROR_NOT_IMPLEMENTED. */
BfCudaStatus bfCudaFlatten(BfMat const *A, BfCudaPlanDesc **desc);
BfCudaStatus bfCudaFlattenEx(BfMat const *A, uint32_t flags, BfCudaPlanDesc **desc);
void bfCudaFlattenFree(BfCudaPlanDesc *desc);

/* Flatten + compile + upload + swap the vtable.  The plan is a snapshot: in-place edits of `A`
 * (bfMatScaleCols, bfMatAddInplace ...) after attaching are not seen; detach, edit, re-attach. */
BfCudaStatus bfCudaMatAttach(BfMat *A, BfCudaAttachOpts const *opts);
/* The same with the operator on SEVERAL GPUs of this process (bf_cuda.h: bfCudaCommInit, bfCudaPlanShard): the
 * reference's callers — bfMatMul in a loop, bfSolveGMRES (src/linalg.c:125,155), the multiple-scattering solve
 * (examples/multiple_scattering/multiple_scattering_context.c:776-794) — then drive all of them from their one
 * host thread with no source change.  mode = BF_CUDA_SHARD_BLOCK_ROWS: every GPU stores 1 / nGpus of the factors
 * (operators larger than one GPU's memory; forward products only: bfMatRmul / bfMatRmulVec raise
 * BF_ERROR_NOT_IMPLEMENTED, bfMatTranspose re-shards the transposed operator at the next product);
 * BF_CUDA_SHARD_RHS_COLS: replicas, the columns of X are split (wide blocks; all directions).
 * `comm` is borrowed and must outlive the attachment.  bfCudaMatGetPlan returns NULL for such an object,
 * bfCudaMatGetShardedPlan the sharded plan; fused permutations are not supported on it. */
BfCudaStatus bfCudaMatAttachSharded(BfMat *A, BfCudaComm *comm, int32_t mode, BfCudaAttachOpts const *opts);
BfCudaShardedPlan *bfCudaMatGetShardedPlan(BfMat const *A);
/* Restore the original vtable and free the plan.  No-op error if `A` is not attached. */
BfCudaStatus bfCudaMatDetach(BfMat *A);
/* The plan behind an attached object (NULL if not attached); borrowed. */
BfCudaPlan *bfCudaMatGetPlan(BfMat const *A);
/* Fuse the caller-side tree-order permutations (bfMatPermuteRows before / after the apply,
 * examples/simple/helm2_bie.c:70-81,139) into an attached object's plan. */
BfCudaStatus bfCudaMatSetPerms(BfMat *A, BfPerm const *inPerm, BfPerm const *outPerm);

/* ---- GPU construction of a fac_helm2 operator's values -------------------------
 * libbfcuda_helm2.so (bf_helm2_record.c) defines the reference's two value builders,
 * bfHelm2GetKernelMatrix and bfHelm2GetReexpansionMatrix (src/helm2.c:282-365), and must come before
 * libbutterfly in the global symbol scope.  Outside a recording they forward to the reference.  Inside
 * one they only note their arguments and return value-less leaves, so the reference's own
 * bfFacHelm2Make / bfFacHelm2MakeMultilevel (src/fac_helm2.c:653-1002) lays out the operator without
 * computing a single Hankel function or SVD:
 *
 *     bfCudaHelm2RecordBegin();
 *     BfMat *A = bfFacHelm2MakeMultilevel(&helm, tree, tree);   // structure only
 *     bfCudaHelm2RecordEnd(&rec);
 *     bfCudaMatAttachHelm2(A, rec, NULL, NULL);                 // values computed on the GPU
 *     bfCudaHelm2RecordingFree(rec);
 *     Y = bfMatMul(A, X);                                       // as before
 *
 * The leaves of `A` hold no host values: after bfCudaMatDetach the object can only be deleted. */
typedef struct BfCudaHelm2Recording BfCudaHelm2Recording;
BfCudaStatus bfCudaHelm2RecordBegin(void);
BfCudaStatus bfCudaHelm2RecordEnd(BfCudaHelm2Recording **rec);
void bfCudaHelm2RecordingFree(BfCudaHelm2Recording *rec);
/* Threshold of the truncated-SVD least squares, as a multiple of the reference's max(m, n) eps sigma_0 + eps
 * (src/mat_dense_complex.c:1796-1806; BfCudaHelm2Build.truncationScale).  0 or 1 = the reference's rule (default);
 * 4 gives an operator 4-13x closer to the dense kernel matrix than the reference's own build. */
BfCudaStatus bfCudaHelm2RecordingSetTruncationScale(BfCudaHelm2Recording *rec, double scale);
/* Map the leaves of a flattened operator (bfCudaFlatten) to the recipes recorded for them: fills
 * `build` (pointers borrowed from `rec`) and sets val = NULL on those leaves. */
BfCudaStatus bfCudaHelm2Resolve(BfCudaHelm2Recording *rec, BfCudaPlanDesc *desc, BfCudaHelm2Build *build);
/* bfCudaMatAttach for an operator laid out inside a recording. */
BfCudaStatus bfCudaMatAttachHelm2(BfMat *A, BfCudaHelm2Recording *rec, BfCudaAttachOpts const *opts,
                                  BfCudaHelm2Stats *stats);

/* Drop-in for bfSolveGMRES (src/linalg.c:47-49, same signature and ownership) on an attached
 * operator: the Arnoldi process runs on the GPU (bfCudaSolveGMRES) instead of round-tripping every
 * Krylov vector through bfMatMul.  One complex right-hand side.  A non-NULL M (left preconditioner) is
 * applied with the reference's own bfMatSolve(M, .) on the host, once per iteration. */
BfMat *bfCudaSolveGMRESMat(BfMat const *A, BfMat const *B, BfMat *X0, BfReal tol, BfSize maxNumIter,
                           BfSize *numIter, BfMat const *M);

#ifdef __cplusplus
}
#endif
#endif /* BF_CUDA_ATTACH_H */
