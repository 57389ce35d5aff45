/* bf_flatten.c — host C: walk a reference `BfMat *` tree once and emit the raw-op list of
 * bf_cuda.h.  Compiled against the reference's headers (-I<reference>/include); its undefined
 * `bf*` symbols are resolved by whichever libbutterfly the host program has loaded.
 *
 * The traversal reproduces, node type by node type, the order in which the reference's own Mul
 * routines visit leaves, so that "list order" == "reference summation order":
 *   MatProduct     last factor first, one fresh intermediate per factor  src/mat_product.c:211-245
 *   MatBlockDiag   k = 0..numBlocks-1                                    src/mat_block_diag.c:387-399
 *   MatBlockCoo    k = 0..numBlocks-1 in storage order                   src/mat_block_coo.c:404-418
 *   MatBlockDense  i outer, j inner, block[i*numColBlocks + j]           src/mat_block_dense.c:534-566
 *   MatSum         terms in order                                        src/mat_sum.c:54
 *   MatCooComplex  the Kapur-Rokhlin correction term bfMatBlockDenseAddInplace wraps into a MatSum
 *                  next to a butterfly (src/mat_block_dense.c:458-510, src/quadrature.c:126-167):
 *                  turned into a few small dense leaves over the clusters of its nonzeros
 *                  (buffers owned by the descriptor)                     src/mat_coo_complex.c:212-262
 */
#include "bf_cuda_attach.h"

#include <stdlib.h>
#include <string.h>

#include <bf/mat_block.h>
#include <bf/mat_block_coo.h>
#include <bf/mat_block_dense.h>
#include <bf/mat_block_diag.h>
#include <bf/mat_coo_complex.h>
#include <bf/mat_dense_complex.h>
#include <bf/mat_dense_real.h>
#include <bf/mat_diag_real.h>
#include <bf/mat_identity.h>
#include <bf/mat_product.h>
#include <bf/mat_sum.h>

typedef struct Flat {
  BfCudaRawOp *ops;
  int64_t numOps, capOps;
  int64_t *bufRows;
  int64_t numBufs, capBufs;
  int dtype; /* -1 until the first dense leaf is seen */
  uint32_t flags; /* BF_CUDA_FLATTEN_* */
  void **owned; /* dense buffers made for sparse leaves */
  int64_t numOwned, capOwned;
  BfCudaStatus status;
} Flat;

/* What bfCudaFlatten hands out: the public descriptor followed by the buffers it owns. */
typedef struct FlatDesc {
  BfCudaPlanDesc desc;
  void **owned;
  int64_t numOwned;
} FlatDesc;

static void *newOwned(Flat *f, size_t bytes) {
  if (f->status != BF_CUDA_OK) return NULL;
  if (f->numOwned == f->capOwned) {
    int64_t cap = f->capOwned ? 2 * f->capOwned : 64;
    void **p = realloc(f->owned, (size_t)cap * sizeof *p);
    if (!p) { f->status = BF_CUDA_ERROR_OUT_OF_MEMORY; return NULL; }
    f->owned = p;
    f->capOwned = cap;
  }
  void *buf = calloc(1, bytes ? bytes : 1);
  if (!buf) { f->status = BF_CUDA_ERROR_OUT_OF_MEMORY; return NULL; }
  f->owned[f->numOwned++] = buf;
  return buf;
}

static void pushOp(Flat *f, BfCudaRawOp const *op) {
  if (f->status != BF_CUDA_OK) return;
  if (f->numOps == f->capOps) {
    int64_t cap = f->capOps ? 2 * f->capOps : 1024;
    BfCudaRawOp *p = realloc(f->ops, (size_t)cap * sizeof *p);
    if (!p) { f->status = BF_CUDA_ERROR_OUT_OF_MEMORY; return; }
    f->ops = p;
    f->capOps = cap;
  }
  f->ops[f->numOps++] = *op;
}

static int64_t newBuf(Flat *f, int64_t rows) {
  if (f->status != BF_CUDA_OK) return -1;
  if (f->numBufs == f->capBufs) {
    int64_t cap = f->capBufs ? 2 * f->capBufs : 256;
    int64_t *p = realloc(f->bufRows, (size_t)cap * sizeof *p);
    if (!p) { f->status = BF_CUDA_ERROR_OUT_OF_MEMORY; return -1; }
    f->bufRows = p;
    f->capBufs = cap;
  }
  f->bufRows[f->numBufs] = rows;
  return f->numBufs++;
}

static void setDtype(Flat *f, int dtype) {
  if (f->dtype < 0) f->dtype = dtype;
  else if (f->dtype != dtype) f->status = BF_CUDA_ERROR_NOT_IMPLEMENTED; /* mixed real/complex */
}

static void emit(Flat *f, BfMat const *mat, int64_t inBuf, int64_t inOff, int64_t outBuf,
                 int64_t outOff);

static void emitBlocks(Flat *f, BfMatBlock const *mb, BfSize numBlocks, BfSize const *rowInd,
                       BfSize const *colInd, int64_t inBuf, int64_t inOff, int64_t outBuf,
                       int64_t outOff) {
  for (BfSize k = 0; k < numBlocks && f->status == BF_CUDA_OK; ++k) {
    BfMat const *block = mb->block[k];
    if (block == NULL) continue;
    BfSize i = rowInd ? rowInd[k] : k, j = colInd ? colInd[k] : k;
    emit(f, block, inBuf, inOff + (int64_t)mb->colOffset[j], outBuf,
         outOff + (int64_t)mb->rowOffset[i]);
  }
}

/* ---- sparse correction terms ------------------------------------------------------------------ */
typedef struct CooEnt {
  int64_t i, j;
  BfComplex z;
} CooEnt;

static int cmpRow(const void *a, const void *b) {
  CooEnt const *x = a, *y = b;
  return x->i != y->i ? (x->i < y->i ? -1 : 1) : (x->j != y->j ? (x->j < y->j ? -1 : 1) : 0);
}

static int cmpCol(const void *a, const void *b) {
  CooEnt const *x = a, *y = b;
  return x->j != y->j ? (x->j < y->j ? -1 : 1) : (x->i != y->i ? (x->i < y->i ? -1 : 1) : 0);
}

/* One dense leaf over the bounding box of e[0..n) if the box is not much larger than its nonzeros,
 * else split at the widest gap of the longer side and recurse. */
static void emitCooCluster(Flat *f, CooEnt *e, int64_t n, BfCudaRawOp const *proto) {
  if (f->status != BF_CUDA_OK || n == 0) return;
  int64_t i0 = e[0].i, i1 = e[0].i, j0 = e[0].j, j1 = e[0].j;
  for (int64_t t = 1; t < n; ++t) {
    if (e[t].i < i0) i0 = e[t].i;
    if (e[t].i > i1) i1 = e[t].i;
    if (e[t].j < j0) j0 = e[t].j;
    if (e[t].j > j1) j1 = e[t].j;
  }
  int64_t h = i1 - i0 + 1, w = j1 - j0 + 1;
  if (h * w <= 8 * n + 64) {
    BfComplex *buf = newOwned(f, (size_t)(h * w) * sizeof(BfComplex));
    if (!buf) return;
    for (int64_t t = 0; t < n; ++t) buf[(e[t].i - i0) * w + (e[t].j - j0)] += e[t].z;
    BfCudaRawOp op = *proto;
    op.kind = BF_CUDA_LEAF_DENSE;
    op.inOff += j0; op.outOff += i0;
    op.m = h; op.k = w; op.ld = w;
    op.val = buf;
    pushOp(f, &op);
    return;
  }
  int byRow = h >= w;
  qsort(e, (size_t)n, sizeof *e, byRow ? cmpRow : cmpCol);
  int64_t cut = n / 2, best = -1;
  for (int64_t t = 1; t < n; ++t) {
    int64_t gap = byRow ? e[t].i - e[t - 1].i : e[t].j - e[t - 1].j;
    if (gap > best) { best = gap; cut = t; }
  }
  emitCooCluster(f, e, cut, proto);
  emitCooCluster(f, e + cut, n - cut, proto);
}

static void emitCooComplex(Flat *f, BfMat const *mat, BfCudaRawOp const *proto) {
  BfMatCooComplex const *coo = bfMatConstToMatCooComplexConst(mat);
  if (mat->props & (BF_MAT_PROPS_TRANS | BF_MAT_PROPS_CONJ)) { f->status = BF_CUDA_ERROR_NOT_IMPLEMENTED; return; }
  setDtype(f, BF_CUDA_COMPLEX_F64);
  int64_t n = (int64_t)coo->numElts, m = (int64_t)bfMatGetNumRows(mat);
  if (n == 0 || f->status != BF_CUDA_OK) return;
  CooEnt *e = malloc((size_t)n * sizeof *e);
  int64_t *lastOfRow = NULL;
  if (!e) { f->status = BF_CUDA_ERROR_OUT_OF_MEMORY; return; }
  int64_t cnt = 0;
  if (f->flags & BF_CUDA_FLATTEN_COO_ACCUMULATE) {
    for (int64_t t = 0; t < n; ++t) e[cnt++] = (CooEnt){(int64_t)coo->rowInd[t], (int64_t)coo->colInd[t], coo->value[t]};
  } else {
    /* The reference's product ASSIGNS row i of the result for every stored element
     * (`*outPtr = *inPtr; *outPtr *= z;`, src/mat_coo_complex.c:248-251): of several elements in one
     * row only the last in storage order survives.  Reproduced here. */
    lastOfRow = malloc((size_t)m * sizeof *lastOfRow);
    if (!lastOfRow) { free(e); f->status = BF_CUDA_ERROR_OUT_OF_MEMORY; return; }
    for (int64_t r = 0; r < m; ++r) lastOfRow[r] = -1;
    for (int64_t t = 0; t < n; ++t) lastOfRow[coo->rowInd[t]] = t;
    for (int64_t r = 0; r < m; ++r)
      if (lastOfRow[r] >= 0) {
        int64_t t = lastOfRow[r];
        e[cnt++] = (CooEnt){r, (int64_t)coo->colInd[t], coo->value[t]};
      }
    free(lastOfRow);
  }
  emitCooCluster(f, e, cnt, proto);
  free(e);
}

static void emit(Flat *f, BfMat const *mat, int64_t inBuf, int64_t inOff, int64_t outBuf,
                 int64_t outOff) {
  if (f->status != BF_CUDA_OK) return;
  BfType type = bfMatGetType(mat);
  BfCudaRawOp op;
  memset(&op, 0, sizeof op);
  op.inBuf = inBuf; op.inOff = inOff; op.outBuf = outBuf; op.outOff = outOff;

  switch (type) {
  case BF_TYPE_MAT_DENSE_COMPLEX: {
    BfMatDenseComplex const *d = bfMatConstToMatDenseComplexConst(mat);
    if (d->colStride != 1) { f->status = BF_CUDA_ERROR_NOT_IMPLEMENTED; return; }
    setDtype(f, BF_CUDA_COMPLEX_F64);
    op.kind = BF_CUDA_LEAF_DENSE;
    op.flags = (uint32_t)mat->props & (BF_CUDA_LEAF_TRANS | BF_CUDA_LEAF_CONJ);
    op.m = (int64_t)bfMatGetNumRows(mat);
    op.k = (int64_t)bfMatGetNumCols(mat);
    op.ld = (int64_t)d->rowStride;
    op.val = d->data;
    if (op.m > 0 && op.k > 0) pushOp(f, &op);
    return;
  }
  case BF_TYPE_MAT_DENSE_REAL: {
    BfMatDenseReal const *d = bfMatConstToMatDenseRealConst(mat);
    if (d->super.colStride != 1) { f->status = BF_CUDA_ERROR_NOT_IMPLEMENTED; return; }
    setDtype(f, BF_CUDA_REAL_F64);
    op.kind = BF_CUDA_LEAF_DENSE;
    op.flags = (uint32_t)mat->props & BF_CUDA_LEAF_TRANS;
    op.m = (int64_t)bfMatGetNumRows(mat);
    op.k = (int64_t)bfMatGetNumCols(mat);
    op.ld = (int64_t)d->super.rowStride;
    op.val = d->data;
    if (op.m > 0 && op.k > 0) pushOp(f, &op);
    return;
  }
  case BF_TYPE_MAT_IDENTITY: {
    op.kind = BF_CUDA_LEAF_IDENTITY;
    op.m = (int64_t)bfMatGetNumRows(mat);
    op.k = (int64_t)bfMatGetNumCols(mat);
    if (op.m > 0 && op.k > 0) pushOp(f, &op);
    return;
  }
  case BF_TYPE_MAT_DIAG_REAL: {
    BfMatDiagReal const *d = bfMatConstToMatDiagRealConst(mat);
    op.kind = BF_CUDA_LEAF_DIAG;
    op.m = (int64_t)bfMatGetNumRows(mat);
    op.k = (int64_t)bfMatGetNumCols(mat);
    op.ld = 1;
    op.val = d->data;
    if (op.m > 0 && op.k > 0) pushOp(f, &op);
    return;
  }
  case BF_TYPE_MAT_ZERO:
    return;
  case BF_TYPE_MAT_COO_COMPLEX:
    emitCooComplex(f, mat, &op);
    return;
  case BF_TYPE_MAT_PRODUCT: {
    BfMatProduct const *prod = bfMatConstToMatProductConst(mat);
    BfSize L = bfMatProductNumFactors(prod);
    if (L == 0) { f->status = BF_CUDA_ERROR_INVALID_ARGUMENT; return; }
    int64_t curBuf = inBuf, curOff = inOff;
    for (BfSize i = L; i-- > 1;) {
      BfMat const *factor = bfMatProductGetFactorConst(prod, i);
      int64_t t = newBuf(f, (int64_t)bfMatGetNumRows(factor));
      emit(f, factor, curBuf, curOff, t, 0);
      curBuf = t; curOff = 0;
      if (f->status != BF_CUDA_OK) return;
    }
    emit(f, bfMatProductGetFactorConst(prod, 0), curBuf, curOff, outBuf, outOff);
    return;
  }
  case BF_TYPE_MAT_SUM: {
    BfMatSum const *sum = (BfMatSum const *)mat;
    BfSize n = bfMatSumGetNumTerms(sum);
    for (BfSize i = 0; i < n; ++i)
      emit(f, bfMatSumGetTermConst(sum, i), inBuf, inOff, outBuf, outOff);
    return;
  }
  case BF_TYPE_MAT_BLOCK_DIAG: {
    BfMatBlock const *mb = bfMatConstToMatBlockConst(mat);
    BfSize n = bfMatBlockDiagNumBlocks(bfMatConstToMatBlockDiagConst(mat));
    emitBlocks(f, mb, n, NULL, NULL, inBuf, inOff, outBuf, outOff);
    return;
  }
  case BF_TYPE_MAT_BLOCK_COO: {
    BfMatBlock const *mb = bfMatConstToMatBlockConst(mat);
    BfMatBlockCoo const *coo = bfMatConstToMatBlockCooConst(mat);
    emitBlocks(f, mb, coo->numBlocks, coo->rowInd, coo->colInd, inBuf, inOff, outBuf, outOff);
    return;
  }
  case BF_TYPE_MAT_BLOCK_DENSE: {
    BfMatBlock const *mb = bfMatConstToMatBlockConst(mat);
    BfSize nr = bfMatBlockGetNumRowBlocks(mb), nc = bfMatBlockGetNumColBlocks(mb);
    for (BfSize i = 0; i < nr; ++i)
      for (BfSize j = 0; j < nc && f->status == BF_CUDA_OK; ++j) {
        BfMat const *block = mb->block[i * nc + j];
        if (block == NULL) continue;
        emit(f, block, inBuf, inOff + (int64_t)mb->colOffset[j], outBuf,
             outOff + (int64_t)mb->rowOffset[i]);
      }
    return;
  }
  default:
    /* MatPerm (no Mul in the reference, src/mat_perm.c:53-61), MatFunc, MatPython, MatCsrReal,
     * MatCooReal (no Mul either), MatDiff, MatGivens: refused. */
    f->status = BF_CUDA_ERROR_NOT_IMPLEMENTED;
    return;
  }
}

static void freeOwned(void **owned, int64_t n) {
  for (int64_t i = 0; i < n; ++i) free(owned[i]);
  free(owned);
}

BfCudaStatus bfCudaFlatten(BfMat const *A, BfCudaPlanDesc **out) { return bfCudaFlattenEx(A, 0, out); }

BfCudaStatus bfCudaFlattenEx(BfMat const *A, uint32_t flags, BfCudaPlanDesc **out) {
  if (!A || !out) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  *out = NULL;
  Flat f;
  memset(&f, 0, sizeof f);
  f.dtype = -1;
  f.flags = flags;
  f.status = BF_CUDA_OK;
  int64_t numRows = (int64_t)bfMatGetNumRows(A), numCols = (int64_t)bfMatGetNumCols(A);
  newBuf(&f, numCols); /* BF_CUDA_BUF_X */
  newBuf(&f, numRows); /* BF_CUDA_BUF_Y */
  emit(&f, A, BF_CUDA_BUF_X, 0, BF_CUDA_BUF_Y, 0);
  if (f.status != BF_CUDA_OK) {
    free(f.ops);
    free(f.bufRows);
    freeOwned(f.owned, f.numOwned);
    return f.status;
  }
  FlatDesc *fd = calloc(1, sizeof *fd);
  if (!fd) { free(f.ops); free(f.bufRows); freeOwned(f.owned, f.numOwned); return BF_CUDA_ERROR_OUT_OF_MEMORY; }
  fd->owned = f.owned;
  fd->numOwned = f.numOwned;
  BfCudaPlanDesc *d = &fd->desc;
  d->abiVersion = BF_CUDA_ABI_VERSION;
  d->dtype = f.dtype < 0 ? BF_CUDA_REAL_F64 : f.dtype;
  d->numRows = numRows;
  d->numCols = numCols;
  d->numBufs = f.numBufs;
  d->bufRows = f.bufRows;
  d->numOps = f.numOps;
  d->ops = f.ops;
  d->device = -1;
  d->buildTranspose = 0;
  *out = d;
  return BF_CUDA_OK;
}

void bfCudaFlattenFree(BfCudaPlanDesc *desc) {
  if (!desc) return;
  FlatDesc *fd = (FlatDesc *)desc; /* desc is the first member */
  free((void *)desc->ops);
  free((void *)desc->bufRows);
  freeOwned(fd->owned, fd->numOwned);
  free(fd);
}
