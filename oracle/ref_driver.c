/* ref_driver.c — thin ctypes-friendly drivers around the UNMODIFIED reference library
 * (oracle/_ref/libbutterfly_ref.so).  TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Modelled on examples/simple/bf_all_blocks.c:96-147 (quadtree -> bfFacHelm2MakeMultilevel ->
 * bfMatMul on a Randn RHS).  Everything numeric is done by the reference; this file only moves
 * pointers across the ctypes boundary.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <bf/fac_helm2.h>
#include <bf/helm2.h>
#include <bf/layer_pot.h>
#include <bf/linalg.h>
#include <bf/mat_dense_complex.h>
#include <bf/mat_dense_real.h>
#include <bf/points.h>
#include <bf/quadtree.h>
#include <bf/rand.h>
#include <bf/util.h>
#include <bf/vec_complex.h>
#include <bf/vec_real.h>

typedef struct RefHelm2 {
  double *pointData;
  BfMatDenseReal *pointsMat;
  BfPoints2 const *points;
  BfQuadtree quadtree;
  BfHelm2 helm;
  BfMat *A;
} RefHelm2;

double refNow(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts); /* not bfToc(): that is CPU time, src/util.c:17-22 */
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

void refSeed(uint64_t seed) { bfSeed(seed); }
void refComplexRandn(int64_t n, double *out) { bfComplexRandn((BfSize)n, (BfComplex *)out); }
void refRealRandn(int64_t n, double *out) { bfRealRandn((BfSize)n, out); }

/* points: n x 2 doubles (copied).  kWave: Helmholtz wavenumber.  Single-layer potential, as in
 * examples/simple/run_bf_all_blocks_test.sh. */
RefHelm2 *refHelm2Build(const double *points, int64_t n, double kWave) {
  RefHelm2 *h = calloc(1, sizeof *h);
  h->pointData = malloc(sizeof(double) * 2 * (size_t)n);
  memcpy(h->pointData, points, sizeof(double) * 2 * (size_t)n);
  h->pointsMat = bfMatDenseRealNewViewFromPtr((BfSize)n, 2, h->pointData, 2, 1);
  h->points = bfPoints2ConstViewFromMatDenseReal(h->pointsMat);
  h->helm.k = kWave;
  h->helm.layerPot = BF_LAYER_POTENTIAL_SINGLE;
  bfQuadtreeInit(&h->quadtree, h->points, NULL);
  h->A = bfFacHelm2MakeMultilevel(&h->helm, &h->quadtree, &h->quadtree);
  return h;
}

BfMat *refHelm2GetMat(RefHelm2 *h) { return h->A; }

/* tree->perm (include/bf/tree.h:37-38) */
void refHelm2GetPerm(RefHelm2 *h, int64_t *perm) {
  BfPerm const *p = bfTreeGetPermConst(bfQuadtreeToTree(&h->quadtree));
  for (BfSize i = 0; i < p->size; ++i) perm[i] = (int64_t)p->index[i];
}

/* Dense kernel matrix in tree order (bf_all_blocks.c:104-123), row-major n x n complex. */
void refHelm2DenseKernel(RefHelm2 *h, double *out) {
  BfTree *tree = bfQuadtreeToTree(&h->quadtree);
  BfPerm *revPerm = bfPermGetReversePerm(bfTreeGetPermConst(tree));
  BfMat *ptsPerm = bfMatCopy(bfMatDenseRealToMat(h->pointsMat));
  bfMatPermuteRows(ptsPerm, revPerm);
  BfPoints2 const *pp = bfPoints2ConstViewFromMat(ptsPerm);
  BfMat *K = bfHelm2GetKernelMatrix(&h->helm, pp, pp, NULL, NULL);
  BfMatDenseComplex *Kd = bfMatToMatDenseComplex(K);
  BfSize n = bfMatGetNumRows(K);
  for (BfSize i = 0; i < n; ++i)
    memcpy(out + 2 * i * n, Kd->data + i * Kd->rowStride, sizeof(BfComplex) * n);
  bfMatDelete(&K);
  bfMatDelete(&ptsPerm);
  bfPermDeinitAndDealloc(&revPerm);
}

void refHelm2Free(RefHelm2 *h) {
  if (!h) return;
  if (h->A) bfMatDelete(&h->A);
  bfQuadtreeDeinit(&h->quadtree);
  free(h->pointData);
  free(h);
}

int64_t refMatNumRows(BfMat const *A) { return (int64_t)bfMatGetNumRows(A); }
int64_t refMatNumCols(BfMat const *A) { return (int64_t)bfMatGetNumCols(A); }
int64_t refMatNumBytes(BfMat const *A) { return (int64_t)bfMatNumBytes(A); }
void refMatTranspose(BfMat *A) { bfMatTranspose(A); }

/* Y = bfMatMul(A, X) through the public vtable (src/mat.c:183): X is n x nrhs row-major complex.
 * Returns seconds spent inside bfMatMul. */
double refMatMulComplex(BfMat const *A, const double *x, int64_t n, int64_t nrhs, double *y) {
  BfMatDenseComplex *X = bfMatDenseComplexNewViewFromPtr((BfSize)n, (BfSize)nrhs, (BfComplex *)x);
  double t0 = refNow();
  BfMat *Y = bfMatMul(A, bfMatDenseComplexToMat(X));
  double t1 = refNow();
  BfMatDenseComplex *Yd = bfMatToMatDenseComplex(Y);
  BfSize m = bfMatGetNumRows(Y);
  for (BfSize i = 0; i < m; ++i)
    memcpy(y + 2 * i * nrhs, Yd->data + i * Yd->rowStride, sizeof(BfComplex) * (size_t)nrhs);
  bfMatDelete(&Y);
  BfMat *Xm = bfMatDenseComplexToMat(X);
  bfMatDelete(&Xm);
  return t1 - t0;
}

/* y = bfMatMulVec(A, x) / bfMatRmulVec(A, x) with BfVecReal (src/mat.c:188,198). */
double refMatMulVecReal(BfMat const *A, const double *x, int64_t n, double *y, int transposed) {
  BfVecReal *xv = bfVecRealNewEmpty((BfSize)n);
  memcpy(xv->data, x, sizeof(double) * (size_t)n);
  double t0 = refNow();
  BfVec *yv = transposed ? bfMatRmulVec(A, bfVecRealToVec(xv)) : bfMatMulVec(A, bfVecRealToVec(xv));
  double t1 = refNow();
  BfVecReal *yr = bfVecToVecReal(yv);
  for (BfSize i = 0; i < yv->size; ++i) y[i] = yr->data[i * yr->stride];
  bfVecDelete(&yv);
  BfVec *xvv = bfVecRealToVec(xv);
  bfVecDelete(&xvv);
  return t1 - t0;
}

/* in-place bfMatPermuteRows on an n x nrhs complex matrix (src/mat_dense_complex.c:741-774) */
void refPermuteRowsComplex(double *x, int64_t n, int64_t nrhs, const int64_t *perm) {
  BfMatDenseComplex *X = bfMatDenseComplexNewViewFromPtr((BfSize)n, (BfSize)nrhs, (BfComplex *)x);
  BfPerm *p = bfPermNewEmpty((BfSize)n);
  for (int64_t i = 0; i < n; ++i) p->index[i] = (BfSize)perm[i];
  bfMatPermuteRows(bfMatDenseComplexToMat(X), p);
  bfPermDeinitAndDealloc(&p);
  BfMat *Xm = bfMatDenseComplexToMat(X);
  bfMatDelete(&Xm);
}

void refReversePerm(const int64_t *perm, int64_t n, int64_t *rev) {
  BfPerm *p = bfPermNewEmpty((BfSize)n);
  for (int64_t i = 0; i < n; ++i) p->index[i] = (BfSize)perm[i];
  BfPerm *r = bfPermGetReversePerm(p);
  for (int64_t i = 0; i < n; ++i) rev[i] = (int64_t)r->index[i];
  bfPermDeinitAndDealloc(&r);
  bfPermDeinitAndDealloc(&p);
}

void refPrintBlocks(BfMat const *A, const char *path) {
  FILE *fp = fopen(path, "w");
  bfPrintBlocks(A, 2, fp);
  fclose(fp);
}

int refOpenblasThreads(int n) {
  extern void openblas_set_num_threads(int);
  extern int openblas_get_num_threads(void);
  if (n > 0) openblas_set_num_threads(n);
  return openblas_get_num_threads();
}
