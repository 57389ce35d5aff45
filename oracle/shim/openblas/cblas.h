/* Prototype-only stand-in for <openblas/cblas.h>, used ONLY to compile the
 * unmodified reference sources for the oracle build (oracle/Makefile).
 * The reference hard-codes `#include <openblas/cblas.h>` in include/bf/blas.h:3-5.
 * These are the standard CBLAS signatures (public Netlib interface); only the
 * routines the reference calls are declared.  Test infrastructure, not product. */
#ifndef ORACLE_SHIM_CBLAS_H
#define ORACLE_SHIM_CBLAS_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 } CBLAS_ORDER;
typedef enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113, CblasConjNoTrans = 114 } CBLAS_TRANSPOSE;
typedef enum CBLAS_UPLO { CblasUpper = 121, CblasLower = 122 } CBLAS_UPLO;
typedef enum CBLAS_DIAG { CblasNonUnit = 131, CblasUnit = 132 } CBLAS_DIAG;
typedef enum CBLAS_SIDE { CblasLeft = 141, CblasRight = 142 } CBLAS_SIDE;
typedef CBLAS_ORDER CBLAS_LAYOUT;
typedef int blasint;

void cblas_zgemm(CBLAS_ORDER, CBLAS_TRANSPOSE, CBLAS_TRANSPOSE, blasint M, blasint N, blasint K,
                 const void *alpha, const void *A, blasint lda, const void *B, blasint ldb,
                 const void *beta, void *C, blasint ldc);
void cblas_dgemm(CBLAS_ORDER, CBLAS_TRANSPOSE, CBLAS_TRANSPOSE, blasint M, blasint N, blasint K,
                 double alpha, const double *A, blasint lda, const double *B, blasint ldb,
                 double beta, double *C, blasint ldc);
void cblas_zgemv(CBLAS_ORDER, CBLAS_TRANSPOSE, blasint M, blasint N, const void *alpha,
                 const void *A, blasint lda, const void *X, blasint incX, const void *beta,
                 void *Y, blasint incY);
void cblas_dgemv(CBLAS_ORDER, CBLAS_TRANSPOSE, blasint M, blasint N, double alpha,
                 const double *A, blasint lda, const double *X, blasint incX, double beta,
                 double *Y, blasint incY);
void cblas_ztrsm(CBLAS_ORDER, CBLAS_SIDE, CBLAS_UPLO, CBLAS_TRANSPOSE, CBLAS_DIAG, blasint M,
                 blasint N, const void *alpha, const void *A, blasint lda, void *B, blasint ldb);
void cblas_ztrsv(CBLAS_ORDER, CBLAS_UPLO, CBLAS_TRANSPOSE, CBLAS_DIAG, blasint N,
                 const void *A, blasint lda, void *X, blasint incX);
void cblas_zdotc_sub(blasint n, const void *x, blasint incx, const void *y, blasint incy, void *ret);
void cblas_zscal(blasint N, const void *alpha, void *X, blasint incX);
void cblas_zdscal(blasint N, double alpha, void *X, blasint incX);
void cblas_dscal(blasint N, double alpha, double *X, blasint incX);
void cblas_daxpy(blasint n, double alpha, const double *x, blasint incx, double *y, blasint incy);
double cblas_dznrm2(blasint N, const void *X, blasint incX);
void openblas_set_num_threads(int);
int openblas_get_num_threads(void);
char *openblas_get_config(void);
#ifdef __cplusplus
}
#endif
#endif
