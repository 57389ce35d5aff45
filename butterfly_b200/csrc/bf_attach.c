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
} Wrapper;
#define WRAPPER_MAGIC 0xBFC0DA01u

static Wrapper *wrapperOf(BfMat const *A) {
  Wrapper *w = (Wrapper *)A->vtbl; /* vt is the first member */
  return w;
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
  BfCudaStatus st = bfCudaApplyHost(w->plan, BF_CUDA_OP_N, Xd->data, (int64_t)Xd->rowStride, (int64_t)nrhs,
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
  BfComplex *xt = malloc(sizeof(BfComplex) * (m * p ? m * p : 1));
  BfComplex *yt = malloc(sizeof(BfComplex) * (n * p ? n * p : 1));
  if (!xt || !yt) {
    free(xt); free(yt);
    bfSetError(BF_ERROR_MEMORY_ERROR);
    return NULL;
  }
  for (BfSize i = 0; i < p; ++i)
    for (BfSize j = 0; j < m; ++j) xt[j * p + i] = Xd->data[i * Xd->rowStride + j];
  BfCudaStatus st = bfCudaApplyHost(w->plan, BF_CUDA_OP_T, xt, (int64_t)p, (int64_t)p, yt, (int64_t)p);
  BfMat *result = NULL;
  if (st == BF_CUDA_OK) {
    BfMatDenseComplex *Y = bfMatDenseComplexNew();
    bfMatDenseComplexInit(Y, p, n);
    for (BfSize i = 0; i < p; ++i)
      for (BfSize j = 0; j < n; ++j) Y->data[i * Y->rowStride + j] = yt[j * p + i];
    result = bfMatDenseComplexToMat(Y);
  } else {
    bfSetError(toBfError(st));
  }
  free(xt); free(yt);
  return result;
}

static BfVec *applyVec(BfMat const *A, BfVec const *x, int op) {
  Wrapper *w = wrapperOf(A);
  BfType type = bfVecGetType(x);
  BfSize nIn = op == BF_CUDA_OP_N ? bfMatGetNumCols(A) : bfMatGetNumRows(A);
  BfSize nOut = op == BF_CUDA_OP_N ? bfMatGetNumRows(A) : bfMatGetNumCols(A);
  if (x->size != nIn) {
    bfSetError(BF_ERROR_INCOMPATIBLE_SHAPES);
    return NULL;
  }
  if (w->dtype == BF_CUDA_REAL_F64 && type == BF_TYPE_VEC_REAL) {
    /* the only operand type the reference's block MulVec accepts, src/mat_block_diag.c:418-428 */
    BfVecReal const *xr = bfVecConstToVecRealConst(x);
    BfVecReal *y = bfVecRealNewEmpty(nOut);
    BfCudaStatus st = bfCudaApplyHost(w->plan, op, xr->data, (int64_t)xr->stride, 1, y->data, (int64_t)y->stride);
    if (st != BF_CUDA_OK) {
      BfVec *yv = bfVecRealToVec(y);
      bfVecDelete(&yv);
      bfSetError(toBfError(st));
      return NULL;
    }
    return bfVecRealToVec(y);
  }
  if (w->dtype == BF_CUDA_COMPLEX_F64 && type == BF_TYPE_VEC_COMPLEX) {
    BfVecComplex const *xc = bfVecConstToVecComplexConst(x);
    BfVecComplex *y = bfVecComplexNew();
    bfVecComplexInit(y, nOut);
    BfCudaStatus st = bfCudaApplyHost(w->plan, op, xc->data, (int64_t)xc->stride, 1, y->data, (int64_t)y->stride);
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

static BfVec *attachedMulVec(BfMat const *A, BfVec const *x) { return applyVec(A, x, BF_CUDA_OP_N); }
static BfVec *attachedRmulVec(BfMat const *A, BfVec const *x) { return applyVec(A, x, BF_CUDA_OP_T); }

static BfSize attachedNumBytes(BfMat const *A) {
  Wrapper *w = wrapperOf(A);
  BfMat *B = (BfMat *)A;
  BfMatVtable *mine = B->vtbl;
  B->vtbl = w->orig;
  BfSize n = bfMatNumBytes(A);
  B->vtbl = mine;
  return n;
}

static void attachedDelete(BfMat **A) {
  Wrapper *w = wrapperOf(*A);
  (*A)->vtbl = w->orig;
  bfCudaPlanDestroy(w->plan);
  w->magic = 0;
  free(w);
  bfMatDelete(A);
}

static int isAttached(BfMat const *A) {
  /* an attached vtable pointer is the address of a heap Wrapper whose Mul slot is ours */
  return A && A->vtbl && A->vtbl->Mul == attachedMul && wrapperOf(A)->magic == WRAPPER_MAGIC;
}

/* bf_helm2_record.c (libbfcuda_helm2.so); absent unless that library is loaded */
extern BfCudaStatus bfCudaHelm2Resolve(BfCudaHelm2Recording *, BfCudaPlanDesc *, BfCudaHelm2Build *) __attribute__((weak));

static BfCudaStatus attachWith(BfMat *A, BfCudaAttachOpts const *opts, BfCudaHelm2Recording *rec,
                               BfCudaHelm2Stats *stats) {
  if (!A) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  if (isAttached(A)) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  if (rec && !bfCudaHelm2Resolve) return BF_CUDA_ERROR_NOT_IMPLEMENTED;
  BfCudaPlanDesc *desc = NULL;
  BfCudaStatus st = bfCudaFlattenEx(A, opts ? opts->flattenFlags : 0u, &desc);
  if (st != BF_CUDA_OK) return st;
  desc->device = opts ? opts->device : -1;
  desc->buildTranspose = opts ? opts->buildTranspose : 0;
  BfCudaHelm2Build build;
  /* recorded leaves: map them to their recipes and strip the (inaccessible) value pointers */
  if (rec) st = bfCudaHelm2Resolve(rec, desc, &build);
  BfCudaPlan *plan = NULL;
  if (st == BF_CUDA_OK) st = bfCudaPlanCreate(desc, &plan);
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
    return BF_CUDA_ERROR_OUT_OF_MEMORY;
  }
  w->vt = *A->vtbl;
  w->orig = A->vtbl;
  w->plan = plan;
  w->dtype = dtype;
  w->magic = WRAPPER_MAGIC;
  w->vt.Mul = attachedMul;
  w->vt.MulVec = attachedMulVec;
  w->vt.Rmul = attachedRmul;
  w->vt.RmulVec = attachedRmulVec;
  /* slots the type leaves empty stay empty (MatSum has neither, src/mat_sum.c:12-17: such an object
   * is freed with its own Deinit after bfCudaMatDetach) */
  if (w->orig->NumBytes) w->vt.NumBytes = attachedNumBytes;
  if (w->orig->Delete) w->vt.Delete = attachedDelete;
  A->vtbl = &w->vt;
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaMatAttach(BfMat *A, BfCudaAttachOpts const *opts) { return attachWith(A, opts, NULL, NULL); }

BfCudaStatus bfCudaMatAttachHelm2(BfMat *A, BfCudaHelm2Recording *rec, BfCudaAttachOpts const *opts,
                                  BfCudaHelm2Stats *stats) {
  if (!rec) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  return attachWith(A, opts, rec, stats);
}

BfCudaStatus bfCudaMatDetach(BfMat *A) {
  if (!isAttached(A)) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  Wrapper *w = wrapperOf(A);
  A->vtbl = w->orig;
  bfCudaPlanDestroy(w->plan);
  w->magic = 0;
  free(w);
  return BF_CUDA_OK;
}

BfCudaPlan *bfCudaMatGetPlan(BfMat const *A) { return isAttached(A) ? wrapperOf(A)->plan : NULL; }

BfCudaStatus bfCudaMatSetPerms(BfMat *A, BfPerm const *inPerm, BfPerm const *outPerm) {
  if (!isAttached(A)) return BF_CUDA_ERROR_INVALID_ARGUMENT;
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
                        ? bfCudaSolveGMRES(wrapperOf(A)->plan, b, x0, tol, (int64_t)maxNumIter, X->data, &it, NULL, NULL)
                        : bfCudaSolveGMRESPrecond(wrapperOf(A)->plan, b, x0, tol, (int64_t)maxNumIter, X->data, &it, NULL,
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
