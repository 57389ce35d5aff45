/* Prototype-only stand-in for <openblas/lapacke.h> (see cblas.h in this directory).
 * Standard LAPACKE C interface; only what the reference calls.  Test infrastructure. */
#ifndef ORACLE_SHIM_LAPACKE_H
#define ORACLE_SHIM_LAPACKE_H
#include <complex.h>
#ifdef __cplusplus
extern "C" {
#endif
#define LAPACK_ROW_MAJOR 101
#define LAPACK_COL_MAJOR 102
#define LAPACK_WORK_MEMORY_ERROR (-1010)
#define LAPACK_TRANSPOSE_MEMORY_ERROR (-1011)
typedef int lapack_int;
typedef double _Complex lapack_complex_double;

lapack_int LAPACKE_zgetrf(int layout, lapack_int m, lapack_int n, lapack_complex_double *a,
                          lapack_int lda, lapack_int *ipiv);
lapack_int LAPACKE_zgetrs(int layout, char trans, lapack_int n, lapack_int nrhs,
                          const lapack_complex_double *a, lapack_int lda, const lapack_int *ipiv,
                          lapack_complex_double *b, lapack_int ldb);
lapack_int LAPACKE_zgesv(int layout, lapack_int n, lapack_int nrhs, lapack_complex_double *a,
                         lapack_int lda, lapack_int *ipiv, lapack_complex_double *b, lapack_int ldb);
lapack_int LAPACKE_zgesvd(int layout, char jobu, char jobvt, lapack_int m, lapack_int n,
                          lapack_complex_double *a, lapack_int lda, double *s,
                          lapack_complex_double *u, lapack_int ldu, lapack_complex_double *vt,
                          lapack_int ldvt, double *superb);
lapack_int LAPACKE_dgesvd(int layout, char jobu, char jobvt, lapack_int m, lapack_int n, double *a,
                          lapack_int lda, double *s, double *u, lapack_int ldu, double *vt,
                          lapack_int ldvt, double *superb);
lapack_int LAPACKE_zlaswp(int layout, lapack_int n, lapack_complex_double *a, lapack_int lda,
                          lapack_int k1, lapack_int k2, const lapack_int *ipiv, lapack_int incx);
#ifdef __cplusplus
}
#endif
#endif
