/* Stand-in for ARPACK-NG's <arpack.h> so that the reference's src/linalg.c compiles
 * (it is needed for bfSolveGMRES, src/linalg.c:47).  The eigen-solvers that call these
 * entry points are NOT part of the oracle; oracle/ref_stubs.c aborts if they are reached. */
#ifndef ORACLE_SHIM_ARPACK_H
#define ORACLE_SHIM_ARPACK_H
#include <stdint.h>
typedef int a_int;
typedef unsigned int a_uint;
void dnaupd_c(a_int *ido, char const *bmat, a_int n, char const *which, a_int nev, double tol,
              double *resid, a_int ncv, double *v, a_int ldv, a_int *iparam, a_int *ipntr,
              double *workd, double *workl, a_int lworkl, a_int *info);
void dneupd_c(a_int rvec, char const *howmny, a_int const *select, double *dr, double *di,
              double *z, a_int ldz, double sigmar, double sigmai, double *workev,
              char const *bmat, a_int n, char const *which, a_int nev, double tol, double *resid,
              a_int ncv, double *v, a_int ldv, a_int *iparam, a_int *ipntr, double *workd,
              double *workl, a_int lworkl, a_int *info);
#endif
