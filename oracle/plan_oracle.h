/* plan_oracle.h — CPU oracle ("port") for the butterfly apply.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this.  The product (libbfcuda.so) never does.
 *
 * It restates, in plain scalar C, what the reference does when it applies a flattened operator:
 * one dense product per leaf, accumulated in the reference's traversal order
 * (src/mat_block_coo.c:404-418, src/mat_block_diag.c:387-399, src/mat_block_dense.c:534-566,
 * src/mat_product.c:225-238; leaf arithmetic = cblas_zgemm / cblas_dgemv semantics,
 * src/mat_dense_complex.c:1754, src/mat_dense_real.c:1358,1428).
 *
 * Pinning: checked against the UNMODIFIED reference's own bfMatMul / bfMatMulVec / bfMatRmulVec
 * (oracle/_ref, built from /root/reference) by tests/test_oracle_vs_reference.py and against the
 * golden vectors in tests/golden/ that oracle/make_golden.py produced with that build.
 */
#ifndef PLAN_ORACLE_H
#define PLAN_ORACLE_H
#include <stdint.h>
#include "../include/bf_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* y (rowsOut x nrhs, row stride ldy) = op(A) x, following desc->ops in order (op N) or in reverse
 * with roles swapped (op T / H).  Returns 0 on success, nonzero on a malformed plan.

 * valsOverride + poolOff[i] (elements), row stride poolLd[i]; used with synthetic pools. */
int planOracleApply(const BfCudaPlanDesc *desc, int32_t op, const void *x, int64_t ldx,
                    int64_t nrhs, void *y, int64_t ldy);

/* E = sum over dense leaves of m*k. */
int64_t planOracleNumEntries(const BfCudaPlanDesc *desc);

/* out[perm[i], :] = in[i, :]  (bfMatDenseComplexPermuteRows, src/mat_dense_complex.c:741-774;
 * bfVecRealPermute, src/vec_real.c:312-330).  elemDoubles = 1 (real) or 2 (complex). */
void planOraclePermuteRows(const int64_t *perm, int64_t n, int64_t nrhs, int32_t elemDoubles,
                           const double *in, int64_t ldin, double *out, int64_t ldout);

#ifdef __cplusplus
}
#endif
#endif
