/* bf_attach.c — host C: route bfMatMul / bfMatMulVec / bfMatRmul / bfMatRmulVec on ONE reference
 * object to its GPU plan by swapping the object's public vtable pointer
 * (include/bf/mat.h:187-188; TODO.md:21-31 anticipates exactly this wrapping).
 *
 * Compiled against the reference's headers; its bf* symbols are resolved by the libbutterfly the
 * host program has loaded.  Results are ordinary reference objects (bfMatDenseComplexNew + Init,
 * bfVecRealNewEmpty: malloc-backed, freed by the caller with bfMatDelete / bfVecDelete,
 * src/mat_dense_complex.c:2164-2187).  Errors follow the reference's convention: bfSetError +
 * NULL (README.md:25-35; note bfSetError asserts in every build, src/error.c:19-23).
 */
#include "bf_cuda_attach.h"

#include <complex.h>
#include <stdlib.h>
#include <string.h>

#include <bf/error.h>
#include <bf/mat_dense_complex.h>
#include <bf/vec_complex.h>
#include <bf/vec_real.h>

typedef struct Wrapper {
  BfMatVtable vt;       /* copy of the original vtable with the apply slots overridden; FIRST */
  BfMatVtable *orig;
  BfCudaPlan *plan;
  int dtype;
  unsigned magic;
  BfCudaAttachOpts opts; /* what the snapshot was taken with (re-used when it is refreshed) */
  int transposed;       /* bfMatTranspose was called an odd number of times since the snapshot: the
                           plan's reverse direction is what bfMatMul means now */
  int dirty;            /* the host object was edited in place since the snapshot (ScaleCols,
                           AddInplace ...): re-flatten before the next apply */
  int valueless;        /* attached through a helm2 recording: the leaves hold no host values, so
                           the object cannot be edited or re-flattened */
  BfCudaShardedPlan *sharded; /* bfCudaMatAttachSharded: the operator lives on several GPUs (plan == NULL then) */
  BfCudaComm *comm;           /* borrowed; must outlive the attachment */
  int shardMode;
  int64_t *inPerm, *outPerm; /* host copies of the fused permutations (survive a refresh) */
  int64_t numInPerm, numOutPerm;
} Wrapper;
#define WRAPPER_MAGIC 0xBFC0DA01u

static Wrapper *wrapperOf(BfMat const *A) {
  Wrapper *w = (Wrapper *)A->vtbl; /* vt is the first member */
  return w;
}

/* Run `stmt` with the object's ORIGINAL vtable in place: the reference's own implementation of a
 * slot sees the object exactly as it was before the attach.  (Needed, not just tidy: GetView / Copy of
 * the block types do `*view = *orig`, src/mat_product.c:31-46 — a view made while the swapped pointer
 * is in place would carry it, and deleting that view would tear down the owner's plan.) */
#define WITH_ORIG_VTABLE(A, w, stmt)      \
  do {                                    \
    BfMat *B_ = (BfMat *)(A);             \
    BfMatVtable *mine_ = B_->vtbl;        \
    B_->vtbl = (w)->orig;                 \
    stmt;                                 \
    B_->vtbl = mine_;                     \
  } while (0)

static void freeWrapper(Wrapper *w) {
  bfCudaPlanDestroy(w->plan);
  bfCudaShardedPlanDestroy(w->sharded);
  w->magic = 0;
  free(w->inPerm);
  free(w->outPerm);
  free(w);
}

/* The plan is a snapshot of the host object.  After an in-place edit the snapshot is retaken the next
 * time the operator is applied, so an attached object never applies stale factors. */
static int refreshIfDirty(BfMat const *A) {
  Wrapper *w = wrapperOf(A);
  if (!w->dirty) return 1;
  BfCudaPlanDesc *desc = NULL;
  BfCudaStatus st = bfCudaFlattenEx(A, w->opts.flattenFlags, &desc);
  BfCudaPlan *plan = NULL;
  if (st == BF_CUDA_OK && w->comm) { /* several GPUs: re-shard the edited operator */
    BfCudaShardedPlan *sp = NULL;
    desc->buildTranspose = w->opts.buildTranspose;
    st = bfCudaPlanShard(desc, w->comm, w->shardMode, NULL, &sp);
    bfCudaFlattenFree(desc);
    if (st != BF_CUDA_OK) {
      bfSetError(BF_ERROR_RUNTIME_ERROR);
      return 0;
    }
    bfCudaShardedPlanDestroy(w->sharded);
    w->sharded = sp;
    w->dirty = 0;
    w->transposed = 0;
    return 1;
  }
  if (st == BF_CUDA_OK) {
    desc->device = w->opts.device;
    desc->buildTranspose = w->opts.buildTranspose;
    st = bfCudaPlanCreate(desc, &plan);
    bfCudaFlattenFree(desc);
  }
  if (st == BF_CUDA_OK && (w->inPerm || w->outPerm)) st = bfCudaPlanSetPerms(plan, w->inPerm, w->outPerm);
  if (st != BF_CUDA_OK) {
    bfCudaPlanDestroy(plan);
    bfSetError(BF_ERROR_RUNTIME_ERROR);
    return 0;
  }
  bfCudaPlanDestroy(w->plan);
  w->plan = plan;
  w->dirty = 0;
  w->transposed = 0; /* the new snapshot is of the object as it stands */
  return 1;
}

/* Direction of the plan that bfMatMul / bfMatMulVec mean right now. */
static int forwardOp(Wrapper const *w) {
  if (!w->transposed) return BF_CUDA_OP_N;
  return w->dtype == BF_CUDA_COMPLEX_F64 ? BF_CUDA_OP_H : BF_CUDA_OP_T; /* src/mat_dense_complex.c:1475-1478 */
}

/* Host-buffer apply on whatever holds the operator: one plan, or a sharded plan over several GPUs. */
static BfCudaStatus applyHostW(Wrapper *w, int op, const void *x, int64_t ldx, int64_t nrhs, void *y, int64_t ldy) {
  if (w->sharded) return bfCudaShardedApplyHost(w->sharded, op, x, ldx, nrhs, y, ldy);
  return bfCudaApplyHost(w->plan, op, x, ldx, nrhs, y, ldy);
}

static enum BfError toBfError(BfCudaStatus st) {
  switch (st) {
  case BF_CUDA_OK: return BF_ERROR_NONE;
  case BF_CUDA_ERROR_INVALID_ARGUMENT: return BF_ERROR_INVALID_ARGUMENTS;
  case BF_CUDA_ERROR_NOT_IMPLEMENTED: return BF_ERROR_NOT_IMPLEMENTED;
  case BF_CUDA_ERROR_OUT_OF_MEMORY: return BF_ERROR_MEMORY_ERROR;
  case BF_CUDA_ERROR_IO: return BF_ERROR_FILE_ERROR;
  case BF_CUDA_ERROR_INCOMPATIBLE_SHAPES: return BF_ERROR_INCOMPATIBLE_SHAPES;
  default: return BF_ERROR_RUNTIME_ERROR;
  }
}

/* Y = A X with X an n x nrhs BfMatDenseComplex (src/mat_block_dense.c:512, mat_product.c:211) */
static BfMat *attachedMul(BfMat const *A, BfMat const *X) {
  if (!refreshIfDirty(A)) return NULL;
  Wrapper *w = wrapperOf(A);
  if (w->dtype != BF_CUDA_COMPLEX_F64 || bfMatGetType(X) != BF_TYPE_MAT_DENSE_COMPLEX) {
    bfSetError(BF_ERROR_NOT_IMPLEMENTED);
    return NULL;
  }
  BfMatDenseComplex const *Xd = bfMatConstToMatDenseComplexConst(X);
  /* transposed views are refused exactly as GetRowRange refuses them, src/mat_dense_complex.c:628 */
  if (Xd->colStride != 1 || (X->props & (BF_MAT_PROPS_TRANS | BF_MAT_PROPS_CONJ))) {
    bfSetError(BF_ERROR_NOT_IMPLEMENTED);
    return NULL;
  }
  BfSize n = bfMatGetNumRows(X), nrhs = bfMatGetNumCols(X);
  if (n != bfMatGetNumCols(A)) {
    bfSetError(BF_ERROR_INCOMPATIBLE_SHAPES);
    return NULL;
  }
  BfMatDenseComplex *Y = bfMatDenseComplexNew();
  bfMatDenseComplexInit(Y, bfMatGetNumRows(A), nrhs);
  BfCudaStatus st = applyHostW(w, forwardOp(w), Xd->data, (int64_t)Xd->rowStride, (int64_t)nrhs,
                                    Y->data, (int64_t)Y->rowStride);
  if (st != BF_CUDA_OK) {
    BfMat *Ym = bfMatDenseComplexToMat(Y);
    bfMatDelete(&Ym);
    bfSetError(toBfError(st));
    return NULL;
  }
  return bfMatDenseComplexToMat(Y);
}

/* X A = (A^T X^T)^T (src/mat_product.c:282-310); X is p x numRows(A) */
static BfMat *attachedRmul(BfMat const *A, BfMat const *X) {
  if (!refreshIfDirty(A)) return NULL;
  Wrapper *w = wrapperOf(A);
  if (w->dtype != BF_CUDA_COMPLEX_F64 || bfMatGetType(X) != BF_TYPE_MAT_DENSE_COMPLEX) {
    bfSetError(BF_ERROR_NOT_IMPLEMENTED);
    return NULL;
  }
  BfMatDenseComplex const *Xd = bfMatConstToMatDenseComplexConst(X);
  if (Xd->colStride != 1 || (X->props & (BF_MAT_PROPS_TRANS | BF_MAT_PROPS_CONJ))) {
    bfSetError(BF_ERROR_NOT_IMPLEMENTED);
    return NULL;
  }
  BfSize p = bfMatGetNumRows(X), m = bfMatGetNumCols(X), n = bfMatGetNumCols(A);
  if (m != bfMatGetNumRows(A)) {
    bfSetError(BF_ERROR_INVALID_ARGUMENTS); /* as src/mat_product.c:287-288 */
    return NULL;
  }
  BfComplex *xt = malloc(sizeof(BfComplex) * (m * p > 0 ? m * p : 1));
  BfComplex *yt = malloc(sizeof(BfComplex) * (n * p > 0 ? n * p : 1));
  if (!xt || !yt) {
    free(xt); free(yt);
    bfSetError(BF_ERROR_MEMORY_ERROR);
    return NULL;
  }
  /* after bfMatTranspose the object is A0^H (A0 = the snapshot), so X A = (conj(A0) X^T)^T =
   * conj(A0 conj(X^T))^T: the forward direction between two conjugations */
  const int cj = w->transposed;
  for (BfSize i = 0; i < p; ++i)
    for (BfSize j = 0; j < m; ++j) {
      BfComplex v = Xd->data[i * Xd->rowStride + j];
      xt[j * p + i] = cj ? conj(v) : v;
    }
  BfCudaStatus st = applyHostW(w, cj ? BF_CUDA_OP_N : BF_CUDA_OP_T, xt, (int64_t)p, (int64_t)p, yt, (int64_t)p);
  BfMat *result = NULL;
  if (st == BF_CUDA_OK) {
    BfMatDenseComplex *Y = bfMatDenseComplexNew();
    bfMatDenseComplexInit(Y, p, n);
    for (BfSize i = 0; i < p; ++i)
      for (BfSize j = 0; j < n; ++j) Y->data[i * Y->rowStride + j] = cj ? conj(yt[j * p + i]) : yt[j * p + i];
    result = bfMatDenseComplexToMat(Y);
  } else {
    bfSetError(toBfError(st));
  }
  free(xt); free(yt);
  return result;
}

/* rmul = 0: y = A x (MulVec); rmul = 1: y = A^T x (RmulVec, src/mat_product.c:312-345). */
static BfVec *applyVec(BfMat const *A, BfVec const *x, int rmul) {
  if (!refreshIfDirty(A)) return NULL;
  Wrapper *w = wrapperOf(A);
  BfType type = bfVecGetType(x);
  BfSize nIn = rmul ? bfMatGetNumRows(A) : bfMatGetNumCols(A);
  BfSize nOut = rmul ? bfMatGetNumCols(A) : bfMatGetNumRows(A);
  if (x->size != nIn) {
    bfSetError(BF_ERROR_INCOMPATIBLE_SHAPES);
    return NULL;
  }
  if (w->dtype == BF_CUDA_REAL_F64 && type == BF_TYPE_VEC_REAL) {
    /* the only operand type the reference's block MulVec accepts, src/mat_block_diag.c:418-428;
     * real: Transpose is the plain transpose, so a transposed object simply swaps the directions */
    int op = (rmul != 0) != (w->transposed != 0) ? BF_CUDA_OP_T : BF_CUDA_OP_N;
    BfVecReal const *xr = bfVecConstToVecRealConst(x);
    BfVecReal *y = bfVecRealNewEmpty(nOut);
    BfCudaStatus st = applyHostW(w, op, xr->data, (int64_t)xr->stride, 1, y->data, (int64_t)y->stride);
    if (st != BF_CUDA_OK) {
      BfVec *yv = bfVecRealToVec(y);
      bfVecDelete(&yv);
      bfSetError(toBfError(st));
      return NULL;
    }
    return bfVecRealToVec(y);
  }
  if (w->dtype == BF_CUDA_COMPLEX_F64 && type == BF_TYPE_VEC_COMPLEX) {
    /* EXTENSION: the reference's block types raise NOT_IMPLEMENTED for a BfVecComplex operand
     * (src/mat_block_diag.c:418-428, src/mat_block_dense.c:586-592), i.e. a reference program cannot
     * reach this branch without aborting; the attached object answers with the product instead
     * (== bfMatMul with a one-column BfMatDenseComplex; pinned by tests/test_gpu_attach_slots.py).
     * A transposed object is A0^H: MulVec = OP_H; RmulVec = (A0^H)^T x = conj(A0 conj(x)). */
    BfVecComplex const *xc = bfVecConstToVecComplexConst(x);
    BfVecComplex *y = bfVecComplexNew();
    bfVecComplexInit(y, nOut);
    int op = rmul ? (w->transposed ? BF_CUDA_OP_N : BF_CUDA_OP_T) : forwardOp(w);
    const int cj = rmul && w->transposed;
    BfComplex *tmp = NULL;
    BfCudaStatus st;
    if (cj) {
      tmp = malloc(sizeof(BfComplex) * (nIn ? nIn : 1));
      for (BfSize i = 0; i < nIn; ++i) tmp[i] = conj(xc->data[i * xc->stride]);
      st = applyHostW(w, op, tmp, 1, 1, y->data, (int64_t)y->stride);
      for (BfSize i = 0; i < nOut; ++i) y->data[i * y->stride] = conj(y->data[i * y->stride]);
      free(tmp);
    } else {
      st = applyHostW(w, op, xc->data, (int64_t)xc->stride, 1, y->data, (int64_t)y->stride);
    }
    if (st != BF_CUDA_OK) {
      BfVec *yv = bfVecComplexToVec(y);
      bfVecDelete(&yv);
      bfSetError(toBfError(st));
      return NULL;
    }
    return bfVecComplexToVec(y);
  }
  bfSetError(BF_ERROR_NOT_IMPLEMENTED);
  return NULL;
}

static BfVec *attachedMulVec(BfMat const *A, BfVec const *x) { return applyVec(A, x, 0); }
static BfVec *attachedRmulVec(BfMat const *A, BfVec const *x) { return applyVec(A, x, 1); }

static BfSize attachedNumBytes(BfMat const *A) {
  Wrapper *w = wrapperOf(A);
  BfSize n = 0;
  WITH_ORIG_VTABLE(A, w, n = bfMatNumBytes(A));
  return n;
}

static void attachedDelete(BfMat **A) {
  Wrapper *w = wrapperOf(*A);
  (*A)->vtbl = w->orig;
  /* A view never owns the plan.  (bfMatGetView / bfMatCopy on an attached object go through the
   * overrides below and hand out objects with the original vtable; a view that carries the swapped
   * pointer can only come from a typed constructor called directly, e.g. bfMatProductGetView.) */
  if (!((*A)->props & BF_MAT_PROPS_VIEW)) freeWrapper(w);
  bfMatDelete(A);
}

/* ---- slots that alias the object: hand out plain reference objects ------------------------------ */
static BfMat *attachedGetView(BfMat *A) {
  BfMat *V = NULL;
  WITH_ORIG_VTABLE(A, wrapperOf(A), V = bfMatGetView(A));
  return V;
}
static BfMat *attachedCopy(BfMat const *A) {
  Wrapper *w = wrapperOf(A);
  if (w->valueless) { /* no host values to copy */
    bfSetError(BF_ERROR_NOT_IMPLEMENTED);
    return NULL;
  }
  BfMat *C = NULL;
  WITH_ORIG_VTABLE(A, w, C = bfMatCopy(A));
  return C;
}

/* ---- slots that edit the object in place ----------------------------------------------------------
 * The edit is made by the reference's own code on the host object; the snapshot is marked stale and
 * retaken by the next apply (refreshIfDirty).  bfMatTranspose is the cheap exception: when the
 * reverse direction was compiled the plan already holds it (reversed factor order + transposed
 * leaves is exactly what bfMatProductTranspose builds, src/mat_product.c:409-424), so only a flag
 * flips.  An operator attached through a helm2 recording has no host values to edit or re-flatten. */
static int beginEdit(BfMat *A) {
  Wrapper *w = wrapperOf(A);
  if (w->valueless) {
    bfSetError(BF_ERROR_NOT_IMPLEMENTED);
    return 0;
  }
  w->dirty = 1;
  return 1;
}
static void attachedTranspose(BfMat *A) {
  Wrapper *w = wrapperOf(A);
  const int need = w->dtype == BF_CUDA_COMPLEX_F64 ? 2 : 1;
  const int hasReverse = (w->opts.buildTranspose & need) && !(w->sharded && w->shardMode == BF_CUDA_SHARD_BLOCK_ROWS);
  if (!w->dirty && hasReverse) {
    w->transposed ^= 1;
  } else if (!beginEdit(A)) {
    return;
  }
  WITH_ORIG_VTABLE(A, w, bfMatTranspose(A));
}
static void attachedScale(BfMat *A, BfComplex c) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatScale(A, c)); }
static void attachedScaleRows(BfMat *A, BfVec const *v) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatScaleRows(A, v)); }
static void attachedScaleCols(BfMat *A, BfVec const *v) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatScaleCols(A, v)); }
static void attachedDivideCols(BfMat *A, BfVec const *v) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatDivideCols(A, v)); }
static void attachedAddInplace(BfMat *A, BfMat const *B) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatAddInplace(A, B)); }
static void attachedSubInplace(BfMat *A, BfMat const *B) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatSubInplace(A, B)); }
static void attachedAddDiag(BfMat *A, BfMat const *B) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatAddDiag(A, B)); }
static void attachedNegate(BfMat *A) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatNegate(A)); }
static void attachedSetRow(BfMat *A, BfSize i, BfVec const *v) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatSetRow(A, i, v)); }
static void attachedSetCol(BfMat *A, BfSize j, BfVec const *v) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatSetCol(A, j, v)); }
static void attachedSetColRange(BfMat *A, BfSize j, BfSize i0, BfSize i1, BfVec const *v) {
  if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatSetColRange(A, j, i0, i1, v));
}
static void attachedSetRowRange(BfMat *A, BfSize i0, BfSize i1, BfMat const *B) {
  if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatSetRowRange(A, i0, i1, B));
}
static void attachedPermuteRows(BfMat *A, BfPerm const *p) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatPermuteRows(A, p)); }
static void attachedPermuteCols(BfMat *A, BfPerm const *p) { if (beginEdit(A)) WITH_ORIG_VTABLE(A, wrapperOf(A), bfMatPermuteCols(A, p)); }

static int isAttached(BfMat const *A) {
  /* an attached vtable pointer is the address of a heap Wrapper whose Mul slot is ours */
  return A && A->vtbl && A->vtbl->Mul == attachedMul && wrapperOf(A)->magic == WRAPPER_MAGIC;
}

/* bf_helm2_record.c (libbfcuda_helm2.so); absent unless that library is loaded */
extern BfCudaStatus bfCudaHelm2Resolve(BfCudaHelm2Recording *, BfCudaPlanDesc *, BfCudaHelm2Build *) __attribute__((weak));

static BfCudaStatus attachWith(BfMat *A, BfCudaAttachOpts const *opts, BfCudaHelm2Recording *rec,
                               BfCudaHelm2Stats *stats, BfCudaComm *comm, int shardMode) {
  if (!A) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  if (isAttached(A)) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  if (rec && !bfCudaHelm2Resolve) return BF_CUDA_ERROR_NOT_IMPLEMENTED;
  if (rec && comm) return BF_CUDA_ERROR_NOT_IMPLEMENTED; /* (a recorded operator is sharded through bfCudaShardDescs*) */
  BfCudaPlanDesc *desc = NULL;
  BfCudaStatus st = bfCudaFlattenEx(A, opts ? opts->flattenFlags : 0u, &desc);
  if (st != BF_CUDA_OK) return st;
  desc->device = opts ? opts->device : -1;
  desc->buildTranspose = opts ? opts->buildTranspose : 0;
  BfCudaHelm2Build build;
  /* recorded leaves: map them to their recipes and strip the (inaccessible) value pointers */
  if (rec) st = bfCudaHelm2Resolve(rec, desc, &build);
  BfCudaPlan *plan = NULL;
  BfCudaShardedPlan *sharded = NULL;
  if (st == BF_CUDA_OK) st = comm ? bfCudaPlanShard(desc, comm, shardMode, NULL, &sharded) : bfCudaPlanCreate(desc, &plan);
  int dtype = desc->dtype;
  bfCudaFlattenFree(desc);
  if (st != BF_CUDA_OK) return st;
  if (rec) {
    st = bfCudaPlanBuildHelm2(plan, &build, stats);
    if (st != BF_CUDA_OK) {
      bfCudaPlanDestroy(plan);
      return st;
    }
  }
  Wrapper *w = calloc(1, sizeof *w);
  if (!w) {
    bfCudaPlanDestroy(plan);
    bfCudaShardedPlanDestroy(sharded);
    return BF_CUDA_ERROR_OUT_OF_MEMORY;
  }
  w->vt = *A->vtbl;
  w->orig = A->vtbl;
  w->plan = plan;
  w->sharded = sharded;
  w->comm = comm;
  w->shardMode = shardMode;
  w->dtype = dtype;
  w->magic = WRAPPER_MAGIC;
  if (opts) w->opts = *opts;
  else w->opts.device = -1;
  w->valueless = rec != NULL;
  w->vt.Mul = attachedMul;
  w->vt.MulVec = attachedMulVec;
  w->vt.Rmul = attachedRmul;
  w->vt.RmulVec = attachedRmulVec;
  /* slots the type leaves empty stay empty (MatSum has neither, src/mat_sum.c:12-17: such an object
   * is freed with its own Deinit after bfCudaMatDetach; calling an empty slot crashes in the
   * reference too) */
#define OVERRIDE(slot, fn) do { if (w->orig->slot) w->vt.slot = fn; } while (0)
  OVERRIDE(NumBytes, attachedNumBytes);
  OVERRIDE(Delete, attachedDelete);
  OVERRIDE(GetView, attachedGetView);
  OVERRIDE(Copy, attachedCopy);
  OVERRIDE(Transpose, attachedTranspose);
  OVERRIDE(Scale, attachedScale);
  OVERRIDE(ScaleRows, attachedScaleRows);
  OVERRIDE(ScaleCols, attachedScaleCols);
  OVERRIDE(DivideCols, attachedDivideCols);
  OVERRIDE(AddInplace, attachedAddInplace);
  OVERRIDE(SubInplace, attachedSubInplace);
  OVERRIDE(AddDiag, attachedAddDiag);
  OVERRIDE(Negate, attachedNegate);
  OVERRIDE(SetRow, attachedSetRow);
  OVERRIDE(SetCol, attachedSetCol);
  OVERRIDE(SetColRange, attachedSetColRange);
  OVERRIDE(SetRowRange, attachedSetRowRange);
  OVERRIDE(PermuteRows, attachedPermuteRows);
  OVERRIDE(PermuteCols, attachedPermuteCols);
#undef OVERRIDE
  A->vtbl = &w->vt;
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaMatAttach(BfMat *A, BfCudaAttachOpts const *opts) { return attachWith(A, opts, NULL, NULL, NULL, 0); }

BfCudaStatus bfCudaMatAttachSharded(BfMat *A, BfCudaComm *comm, int32_t mode, BfCudaAttachOpts const *opts) {
  if (!comm) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  return attachWith(A, opts, NULL, NULL, comm, mode);
}

BfCudaShardedPlan *bfCudaMatGetShardedPlan(BfMat const *A) { return isAttached(A) ? wrapperOf(A)->sharded : NULL; }

BfCudaStatus bfCudaMatAttachHelm2(BfMat *A, BfCudaHelm2Recording *rec, BfCudaAttachOpts const *opts,
                                  BfCudaHelm2Stats *stats) {
  if (!rec) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  return attachWith(A, opts, rec, stats, NULL, 0);
}

BfCudaStatus bfCudaMatDetach(BfMat *A) {
  if (!isAttached(A)) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  Wrapper *w = wrapperOf(A);
  A->vtbl = w->orig;
  freeWrapper(w);
  return BF_CUDA_OK;
}

/* The plan as a snapshot of the object as it stands NOW: refreshed after in-place edits, and re-taken
 * (forward direction = the transposed object) if bfMatTranspose flipped it. */
static BfCudaPlan *currentForwardPlan(BfMat const *A) {
  Wrapper *w = wrapperOf(A);
  if (w->sharded) { /* no single plan behind a sharded attachment */
    bfSetError(BF_ERROR_NOT_IMPLEMENTED);
    return NULL;
  }
  if (w->transposed) {
    if (w->valueless) {
      bfSetError(BF_ERROR_NOT_IMPLEMENTED);
      return NULL;
    }
    w->dirty = 1;
  }
  return refreshIfDirty(A) ? w->plan : NULL;
}

BfCudaPlan *bfCudaMatGetPlan(BfMat const *A) { return isAttached(A) ? currentForwardPlan(A) : NULL; }

BfCudaStatus bfCudaMatSetPerms(BfMat *A, BfPerm const *inPerm, BfPerm const *outPerm) {
  if (!isAttached(A)) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  if (wrapperOf(A)->sharded) return BF_CUDA_ERROR_NOT_IMPLEMENTED;
  int64_t *in = NULL, *out = NULL;
  if (inPerm) {
    in = malloc(sizeof(int64_t) * (inPerm->size ? inPerm->size : 1));
    for (BfSize i = 0; i < inPerm->size; ++i) in[i] = (int64_t)inPerm->index[i];
  }
  if (outPerm) {
    out = malloc(sizeof(int64_t) * (outPerm->size ? outPerm->size : 1));
    for (BfSize i = 0; i < outPerm->size; ++i) out[i] = (int64_t)outPerm->index[i];
  }
  BfCudaStatus st = BF_CUDA_OK;
  if ((inPerm && inPerm->size != bfMatGetNumCols(A)) || (outPerm && outPerm->size != bfMatGetNumRows(A)))
    st = BF_CUDA_ERROR_INCOMPATIBLE_SHAPES;
  else
    st = bfCudaPlanSetPerms(wrapperOf(A)->plan, in, out);
  if (st == BF_CUDA_OK) { /* kept for refreshIfDirty */
    Wrapper *w = wrapperOf(A);
    free(w->inPerm);
    free(w->outPerm);
    w->inPerm = in;
    w->outPerm = out;
    return st;
  }
  free(in);
  free(out);
  return st;
}

/* bfSolveGMRES (src/linalg.c:47) with the Krylov basis resident on the GPU.  Same signature, same
 * ownership (a non-NULL X0 is consumed: the reference deletes it before returning,
 * src/linalg.c:312-313), same *numIter.  A must be attached, B (and X0) n x 1 BfMatDenseComplex with
 * unit column stride.  A non-NULL M is the reference's left preconditioner: bfMatSolve(M, .) runs on
 * the host exactly as in the reference (src/linalg.c:131,159 — block LU solves in the examples), the
 * vector crossing PCIe once each way per iteration.  Anything else raises BF_ERROR_NOT_IMPLEMENTED like
 * an unimplemented vtable slot would. */
static int solveWithM(void *ctx, const void *in, void *out, int64_t n) {
  BfMat const *M = ctx;
  BfMatDenseComplex *W = bfMatDenseComplexNewViewFromPtr((BfSize)n, 1, (BfComplex *)in);
  BfMat *Wm = bfMatDenseComplexToMat(W);
  BfMat *Z = bfMatSolve(M, Wm);
  bfMatDelete(&Wm);
  if (Z == NULL) return 1;
  int bad = bfMatGetType(Z) != BF_TYPE_MAT_DENSE_COMPLEX || bfMatGetNumRows(Z) != (BfSize)n || bfMatGetNumCols(Z) != 1;
  if (!bad) {
    BfMatDenseComplex const *Zd = bfMatConstToMatDenseComplexConst(Z);
    for (int64_t i = 0; i < n; ++i) ((BfComplex *)out)[i] = Zd->data[(BfSize)i * Zd->rowStride];
  }
  bfMatDelete(&Z);
  return bad;
}

BfMat *bfCudaSolveGMRESMat(BfMat const *A, BfMat const *B, BfMat *X0, BfReal tol, BfSize maxNumIter,
                           BfSize *numIter, BfMat const *M) {
  if (!isAttached(A) || bfMatGetType(B) != BF_TYPE_MAT_DENSE_COMPLEX ||
      (X0 != NULL && bfMatGetType(X0) != BF_TYPE_MAT_DENSE_COMPLEX) || bfMatGetNumCols(B) != 1) {
    bfSetError(BF_ERROR_NOT_IMPLEMENTED);
    return NULL;
  }
  if (maxNumIter == 0 || bfMatGetNumRows(A) != bfMatGetNumCols(A) || bfMatGetNumRows(B) != bfMatGetNumRows(A) ||
      (X0 != NULL && (bfMatGetNumRows(X0) != bfMatGetNumRows(A) || bfMatGetNumCols(X0) != 1)) ||
      (M != NULL && (bfMatGetNumRows(M) != bfMatGetNumRows(A) || bfMatGetNumCols(M) != bfMatGetNumRows(A)))) {
    bfSetError(BF_ERROR_INVALID_ARGUMENTS); /* src/linalg.c:80-110 */
    return NULL;
  }
  BfSize n = bfMatGetNumRows(A);
  BfCudaPlan *plan = currentForwardPlan(A);
  if (plan == NULL) return NULL;
  BfMatDenseComplex const *Bd = bfMatConstToMatDenseComplexConst(B);
  BfMatDenseComplex const *X0d = X0 ? bfMatConstToMatDenseComplexConst(X0) : NULL;
  BfComplex *b = malloc(sizeof(BfComplex) * (n ? n : 1)), *x0 = NULL;
  for (BfSize i = 0; i < n; ++i) b[i] = Bd->data[i * Bd->rowStride];
  if (X0d) {
    x0 = malloc(sizeof(BfComplex) * (n ? n : 1));
    for (BfSize i = 0; i < n; ++i) x0[i] = X0d->data[i * X0d->rowStride];
  }
  BfMatDenseComplex *X = bfMatDenseComplexNew();
  bfMatDenseComplexInit(X, n, 1);
  int64_t it = 0;
  BfCudaStatus st = M == NULL
                        ? bfCudaSolveGMRES(plan, b, x0, tol, (int64_t)maxNumIter, X->data, &it, NULL, NULL)
                        : bfCudaSolveGMRESPrecond(plan, b, x0, tol, (int64_t)maxNumIter, X->data, &it, NULL,
                                                  NULL, solveWithM, (void *)M);
  free(b);
  free(x0);
  if (X0 != NULL) bfMatDelete(&X0);
  BfMat *result = bfMatDenseComplexToMat(X);
  if (st != BF_CUDA_OK) {
    bfMatDelete(&result);
    bfSetError(toBfError(st));
    return NULL;
  }
  if (numIter) *numIter = (BfSize)it;
  return result;
}
