/* plan_oracle.c — CPU oracle ("port") of the butterfly apply over a flattened operator.
 * TEST INFRASTRUCTURE, NOT PRODUCT: see plan_oracle.h for who may use it and how it is pinned.
 *
 * Scalar loops only, no BLAS, one thread: every raw op is applied exactly as the reference
 * applies the corresponding leaf and the partial results are summed in list order.
 */
#include "plan_oracle.h"

#include <stdlib.h>
#include <string.h>

typedef struct {
  double *data; /* rows x nrhs x elemDoubles, dense */
  int64_t rows;
} OracleBuf;

static int leafActsConjTrans(const BfCudaRawOp *op, int complexDtype) {
  /* getCblasTranspose, src/mat_dense_complex.c:27-35: either bit => CblasConjTrans.
   * Real leaves: TRANS => CblasTrans (src/mat_dense_real.c:1345-1358). */
  if (complexDtype) return (op->flags & (BF_CUDA_LEAF_TRANS | BF_CUDA_LEAF_CONJ)) != 0;
  return (op->flags & BF_CUDA_LEAF_TRANS) != 0;
}

/* dst[dOff + i, :] += M(i, j) * src[sOff + j, :], M = the leaf's acting matrix (rowsOut x rowsIn),
 * optionally replaced by its transpose / adjoint (applyT / applyH) for the T / H directions. */
static void applyDense(const BfCudaRawOp *op, int cplx, int dirT, int dirH, const double *src,
                       int64_t sOff, double *dst, int64_t dOff, int64_t nrhs) {
  const int storedT = leafActsConjTrans(op, cplx); /* acting = stored^T (conj if complex) */
  const int e = cplx ? 2 : 1;
  /* acting matrix Aact is m x k.  direction N: out(m) += Aact in(k).
   * direction T: in-role(k) += Aact^T out-role(m).  H: += Aact^H. */
  const int64_t rowsOut = (dirT || dirH) ? op->k : op->m;
  const int64_t rowsIn = (dirT || dirH) ? op->m : op->k;
  const double *S = (const double *)op->val;
  for (int64_t i = 0; i < rowsOut; ++i) {
    double *drow = dst + (dOff + i) * nrhs * e;
    for (int64_t j = 0; j < rowsIn; ++j) {
      /* element (i, j) of the matrix being applied = Aact(i,j) for N, Aact(j,i) (conj for H)
       * for T/H.  Aact(p,q) = stored(p,q) or conj(stored(q,p)). */
      int64_t p = (dirT || dirH) ? j : i, q = (dirT || dirH) ? i : j;
      int conj = dirH;
      int64_t sr, sc;
      if (storedT) { sr = q; sc = p; conj ^= cplx; } else { sr = p; sc = q; }
      const double *sv = S + (sr * op->ld + sc) * e;
      const double *srow = src + (sOff + j) * nrhs * e;
      if (!cplx) {
        const double a = sv[0];
        for (int64_t c = 0; c < nrhs; ++c) drow[c] += a * srow[c];
      } else {
        const double ar = sv[0], ai = conj ? -sv[1] : sv[1];
        for (int64_t c = 0; c < nrhs; ++c) {
          const double xr = srow[2 * c], xi = srow[2 * c + 1];
          drow[2 * c] += ar * xr - ai * xi;
          drow[2 * c + 1] += ar * xi + ai * xr;
        }
      }
    }
  }
}

static void applyIdentityOrDiag(const BfCudaRawOp *op, int cplx, int dirTH, const double *src,
                                int64_t sOff, double *dst, int64_t dOff, int64_t nrhs) {
  (void)dirTH; /* identity and real diagonal are symmetric and real: N == T == H */
  const int e = cplx ? 2 : 1;
  const int64_t n = op->m < op->k ? op->m : op->k;
  const double *d = (const double *)op->val;
  for (int64_t i = 0; i < n; ++i) {
    const double s = (op->kind == BF_CUDA_LEAF_DIAG) ? d[i] : 1.0;
    double *drow = dst + (dOff + i) * nrhs * e;
    const double *srow = src + (sOff + i) * nrhs * e;
    for (int64_t c = 0; c < nrhs * e; ++c) drow[c] += s * srow[c];
  }
}

int64_t planOracleNumEntries(const BfCudaPlanDesc *desc) {
  int64_t E = 0;
  for (int64_t i = 0; i < desc->numOps; ++i)
    if (desc->ops[i].kind == BF_CUDA_LEAF_DENSE) E += desc->ops[i].m * desc->ops[i].k;
  return E;
}

int planOracleApply(const BfCudaPlanDesc *desc, int32_t op, const void *x, int64_t ldx,
                    int64_t nrhs, void *y, int64_t ldy) {
  if (!desc || desc->numBufs < 2 || nrhs < 0) return 1;
  const int cplx = desc->dtype == BF_CUDA_COMPLEX_F64;
  const int e = cplx ? 2 : 1;
  const int dirT = op == BF_CUDA_OP_T, dirH = op == BF_CUDA_OP_H;
  const int rev = dirT || dirH;
  if (dirH && !cplx) return 1;
  if (desc->bufRows[BF_CUDA_BUF_X] != desc->numCols || desc->bufRows[BF_CUDA_BUF_Y] != desc->numRows)
    return 2;

  OracleBuf *buf = calloc((size_t)desc->numBufs, sizeof *buf);
  if (!buf) return 3;
  int rc = 0;
  for (int64_t b = 0; b < desc->numBufs; ++b) {
    buf[b].rows = desc->bufRows[b];
    size_t n = (size_t)buf[b].rows * (size_t)nrhs * e;
    buf[b].data = calloc(n ? n : 1, sizeof(double)); /* bfMatZerosLike */
    if (!buf[b].data) { rc = 3; goto done; }
  }
  /* load the operand into the buffer that plays "input" in this direction */
  {
    const int64_t bIn = rev ? BF_CUDA_BUF_Y : BF_CUDA_BUF_X;
    const double *xs = (const double *)x;
    for (int64_t i = 0; i < buf[bIn].rows; ++i)
      memcpy(buf[bIn].data + i * nrhs * e, xs + i * ldx * e, sizeof(double) * nrhs * e);
  }
  for (int64_t n = 0; n < desc->numOps; ++n) {
    const BfCudaRawOp *o = &desc->ops[rev ? desc->numOps - 1 - n : n];
    if (o->inBuf < 0 || o->inBuf >= desc->numBufs || o->outBuf < 0 || o->outBuf >= desc->numBufs ||
        o->m < 0 || o->k < 0 || o->inOff < 0 || o->outOff < 0 ||
        o->inOff + o->k > buf[o->inBuf].rows || o->outOff + o->m > buf[o->outBuf].rows) {
      rc = 4;
      goto done;
    }
    const int64_t sB = rev ? o->outBuf : o->inBuf, sOff = rev ? o->outOff : o->inOff;
    const int64_t dB = rev ? o->inBuf : o->outBuf, dOff = rev ? o->inOff : o->outOff;
    if (o->kind == BF_CUDA_LEAF_DENSE) {
      if (!o->val) { rc = 5; goto done; }
      applyDense(o, cplx, dirT, dirH, buf[sB].data, sOff, buf[dB].data, dOff, nrhs);
    } else if (o->kind == BF_CUDA_LEAF_IDENTITY || o->kind == BF_CUDA_LEAF_DIAG) {
      if (o->kind == BF_CUDA_LEAF_DIAG && !o->val) { rc = 5; goto done; }
      applyIdentityOrDiag(o, cplx, rev, buf[sB].data, sOff, buf[dB].data, dOff, nrhs);
    } else {
      rc = 6;
      goto done;
    }
  }
  {
    const int64_t bOut = rev ? BF_CUDA_BUF_X : BF_CUDA_BUF_Y;
    double *ys = (double *)y;
    for (int64_t i = 0; i < buf[bOut].rows; ++i)
      memcpy(ys + i * ldy * e, buf[bOut].data + i * nrhs * e, sizeof(double) * nrhs * e);
  }
done:
  for (int64_t b = 0; b < desc->numBufs; ++b) free(buf[b].data);
  free(buf);
  return rc;
}

void planOraclePermuteRows(const int64_t *perm, int64_t n, int64_t nrhs, int32_t elemDoubles,
                           const double *in, int64_t ldin, double *out, int64_t ldout) {
  for (int64_t i = 0; i < n; ++i)
    memcpy(out + perm[i] * ldout * elemDoubles, in + i * ldin * elemDoubles,
           sizeof(double) * nrhs * elemDoubles);
}
