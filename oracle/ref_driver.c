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
#include <bf/lu.h>
#include <bf/mat_block_dense.h>
#include <bf/mat_block_diag.h>
#include <bf/ptr_array.h>
#include <complex.h>
#include <bf/mat_coo_complex.h>
#include <bf/mat_dense_complex.h>
#include <bf/mat_dense_real.h>
#include <bf/mat_sum.h>
#include <bf/points.h>
#include <bf/quadtree.h>
#include <bf/rand.h>
#include <bf/util.h>
#include <bf/vec_complex.h>
#include <bf/vec_real.h>
#include <bf/vectors.h>

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
  if (h->A && bfMatGetType(h->A) == BF_TYPE_MAT_SUM) { /* MatSum has no Delete slot (src/mat_sum.c:12-17) */
    BfMatSum *sum = (BfMatSum *)h->A;
    bfMatSumDeinitAndDealloc(&sum);
    h->A = NULL;
  }
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

/* ---- streamed (real, LBO-style) butterfly ----------------------------------------------------
 * Restates the 40 lines of bfLboFeedFacStreamerNextEigenband (src/lbo.c:96-137) with the
 * eigenband supplied by the caller instead of bfGetEigenband (ARPACK/UMFPACK are absent here,
 * SURVEY.md §8c): octree row tree on the points (examples/covariance/lbo_cov.c:88-96), empty
 * frequency interval tree of valence 2 (lbo_cov.c:113-116), BfFacSpec as lbo_cov.c:121-129,
 * then per column leaf: insert the band's frequencies, permute the band's rows into row-tree
 * order, bfFacStreamerFeed.  The result is bfFacSpanGetMat (lbo_cov.c:189-190). */
#include <bf/fac_span.h>
#include <bf/fac_streamer.h>
#include <bf/interval_tree.h>
#include <bf/interval_tree_node.h>
#include <bf/octree.h>
#include <bf/tree.h>
#include <bf/tree_node.h>

typedef struct RefStream {
  BfPoints3 *points;
  BfOctree octree;
  BfTree *rowTree;
  BfPerm *revRowPerm;
  BfIntervalTree *freqTree;
  BfPoints1 *freqs;
  BfFacStreamer *streamer;
  BfMat *Phi; /* bfFacSpanGetMat view, created on demand */
  BfFacSpan *span;
  int64_t n;
} RefStream;

RefStream *refStreamBegin(const double *xyz, int64_t n, double freqMax, int64_t freqTreeDepth, double tol,
                          int64_t rowTreeOffset, int64_t minNumRows, int64_t minNumCols) {
  RefStream *s = calloc(1, sizeof *s);
  s->n = n;
  s->points = bfPoints3NewWithDefaultCapacity();
  for (int64_t i = 0; i < n; ++i) {
    BfPoint3 p = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    bfPoints3Append(s->points, p);
  }
  bfOctreeInit(&s->octree, s->points, NULL, /* maxLeafSize: */ 1);
  s->rowTree = bfOctreeToTree(&s->octree);
  s->revRowPerm = bfPermGetReversePerm(bfTreeGetPermConst(s->rowTree));
  if (freqTreeDepth < 0) freqTreeDepth = (int64_t)bfTreeGetMaxDepth(s->rowTree) - 3;
  s->freqTree = bfIntervalTreeNew();
  bfIntervalTreeInitEmpty(s->freqTree, 0, freqMax, 2, (BfSize)freqTreeDepth);
  s->freqs = bfPoints1New();
  bfPoints1InitEmpty(s->freqs, BF_ARRAY_DEFAULT_CAPACITY);
  BfFacSpec spec = {
    .rowTree = s->rowTree,
    .colTree = bfIntervalTreeToTree(s->freqTree),
    .rowTreeInitDepth = (BfSize)rowTreeOffset,
    .colTreeInitDepth = (BfSize)freqTreeDepth,
    .tol = tol,
    .minNumRows = (BfSize)minNumRows,
    .minNumCols = (BfSize)minNumCols,
    .compareRelativeErrors = false,
  };
  s->streamer = bfFacStreamerNew();
  bfFacStreamerInit(s->streamer, &spec);
  return s;
}

int64_t refStreamRowTreeDepth(RefStream *s) { return (int64_t)bfTreeGetMaxDepth(s->rowTree); }
void refStreamGetRowPerm(RefStream *s, int64_t *perm) {
  BfPerm const *p = bfTreeGetPermConst(s->rowTree);
  for (BfSize i = 0; i < p->size; ++i) perm[i] = (int64_t)p->index[i];
}
int refStreamIsDone(RefStream *s) { return bfFacStreamerIsDone(s->streamer) ? 1 : 0; }

/* The frequency interval [a, b) of the column leaf the streamer expects next (src/lbo.c:40-66). */
int refStreamCurrentBand(RefStream *s, double *a, double *b, int *leftmost, int *rightmost) {
  BfTreeNode *node = bfFacStreamerGetCurrentColumnNode(s->streamer);
  if (!node || !bfTreeNodeIsLeaf(node)) return 1;
  BfIntervalTreeNode const *in = bfTreeNodeConstToIntervalTreeNodeConst(node);
  *a = in->a; *b = in->b; *leftmost = in->isLeftmost; *rightmost = in->isRightmost;
  return 0;
}

/* phi: n x m row-major band in ORIGINAL point order; freqs: its m frequencies, ascending. */
int refStreamFeed(RefStream *s, const double *phi, int64_t m, const double *freqs) {
  BfTreeNode *node = bfFacStreamerGetCurrentColumnNode(s->streamer);
  BfPoints1 *newFreqs = bfPoints1New();
  bfPoints1InitEmpty(newFreqs, BF_ARRAY_DEFAULT_CAPACITY);
  for (int64_t j = 0; j < m; ++j) bfPoints1Append(newFreqs, freqs[j]);
  bfPoints1InsertPointsSorted(s->freqs, newFreqs);
  bfIntervalTreeSetPoints(s->freqTree, s->freqs, /* rebuildTree: */ false);
  if ((int64_t)bfTreeNodeGetNumPoints(node) != m) { bfPoints1Delete(&newFreqs); return 2; } /* src/lbo.c:126 */
  BfMatDenseReal *Phi = bfMatDenseRealNewZeros((BfSize)s->n, (BfSize)m);
  for (int64_t i = 0; i < s->n; ++i)
    memcpy(Phi->data + i * Phi->super.rowStride, phi + i * m, sizeof(double) * (size_t)m);
  BfMat *PhiMat = bfMatDenseRealToMat(Phi);
  bfMatPermuteRows(PhiMat, s->revRowPerm);
  bfFacStreamerFeed(s->streamer, PhiMat);
  bfMatDelete(&PhiMat);
  bfPoints1Delete(&newFreqs);
  return 0;
}

BfMat *refStreamGetMat(RefStream *s) {
  if (!s->Phi) {
    s->span = bfFacStreamerGetFacSpan(s->streamer);
    s->Phi = bfFacSpanGetMat(s->span, BF_POLICY_VIEW);
  }
  return s->Phi;
}
int64_t refStreamNumFreqs(RefStream *s) { return (int64_t)s->freqs->size; }

/* ---- second-kind system + GMRES (examples/multiple_scattering/multiple_scattering_context.c:456-487) */
#include <bf/mat_diag_real.h>

/* K <- K diag(w) + I/2 on the multilevel butterfly: bfMatScaleCols with constant trapezoid weights
 * and bfMatAddInplace with bfMatDiagRealNewConstant(n, n, 0.5), exactly as the example does. */
void refHelm2MakeSecondKind(RefHelm2 *h, double weight) {
  BfSize n = bfMatGetNumRows(h->A);
  BfVecReal *w = bfVecRealNewEmpty(n);
  for (BfSize i = 0; i < n; ++i) w->data[i] = weight;
  bfMatScaleCols(h->A, bfVecRealToVec(w));
  BfVec *wv = bfVecRealToVec(w);
  bfVecDelete(&wv);
  BfMat *oneHalfEye = bfMatDiagRealToMat(bfMatDiagRealNewConstant(n, n, 0.5));
  bfMatAddInplace(h->A, oneHalfEye);
  bfMatDelete(&oneHalfEye);
}

/* Kapur-Rokhlin correction of the multilevel butterfly as multiple_scattering_context.c:467-472 applies
 * it (one closed curve: bfHelm2ApplyKrCorrectionTree, src/helm2.c:398-414).  Butterflied blocks that
 * receive correction entries become MatSum [MatProduct, MatCooComplex] (src/mat_block_dense.c:493-499).
 * The single-layer kernel ignores the normals; a zero array keeps krComplexKernel's indexing valid. */
void refHelm2ApplyKrCorrection(RefHelm2 *h, int64_t order) {
  BfSize n = bfMatGetNumRows(h->A);
  BfVectors2 normals = {.data = calloc(n, sizeof(BfVector2)), .size = n, .capacity = n, .isView = true};
  bfHelm2ApplyKrCorrectionTree(&h->helm, (BfSize)order, h->points, &normals, bfQuadtreeToTree(&h->quadtree), h->A);
  free(normals.data);
}

/* A <- MatSum [A, C] for an n x n MatCooComplex C (tree order): the node bfMatBlockDenseAddInplace
 * builds around a butterflied block that receives correction entries (src/mat_block_dense.c:493-499),
 * here with terms that stay alive — the reference deletes the sparse term it has just put into the
 * sum (`next: bfMatDelete(&otherBlock)`, :502-503), so sums made by that routine cannot be applied. */
void refHelm2SumWithCoo(RefHelm2 *h, int64_t nnz, const int64_t *rowInd, const int64_t *colInd, const double *values) {
  BfSize n = bfMatGetNumRows(h->A);
  BfMatCooComplex *coo = bfMatCooComplexNew();
  bfMatCooComplexInitEmpty(coo, n, n, (BfSize)nnz);
  for (int64_t t = 0; t < nnz; ++t) {
    coo->rowInd[t] = (BfSize)rowInd[t];
    coo->colInd[t] = (BfSize)colInd[t];
    coo->value[t] = values[2 * t] + I * values[2 * t + 1];
  }
  BfMatSum *sum = bfMatSumNew();
  bfMatSumInit(sum);
  bfMatSumAddTerm(sum, h->A);
  bfMatSumAddTerm(sum, bfMatCooComplexToMat(coo));
  h->A = bfMatSumToMat(sum);
}

/* X = bfSolveGMRES(A, B, NULL, tol, maxNumIter, &numIter, NULL) (src/linalg.c:47), B n x 1 complex.
 * Returns seconds; writes the iteration count. */
double refSolveGMRES(BfMat const *A, const double *b, int64_t n, double tol, int64_t maxNumIter, double *x, int64_t *numIter) {
  BfMatDenseComplex *B = bfMatDenseComplexNewViewFromPtr((BfSize)n, 1, (BfComplex *)b);
  BfSize it = 0;
  double t0 = refNow();
  BfMat *X = bfSolveGMRES(A, bfMatDenseComplexToMat(B), NULL, tol, (BfSize)maxNumIter, &it, NULL);
  double t1 = refNow();
  BfMatDenseComplex *Xd = bfMatToMatDenseComplex(X);
  for (int64_t i = 0; i < n; ++i) memcpy(x + 2 * i, Xd->data + i * Xd->rowStride, sizeof(BfComplex));
  bfMatDelete(&X);
  BfMat *Bm = bfMatDenseComplexToMat(B);
  bfMatDelete(&Bm);
  *numIter = (int64_t)it;
  return t1 - t0;
}

/* Helpers for calling a bfSolveGMRES-shaped function pointer (the reference's own, or the attach
 * layer's bfCudaSolveGMRESMat) from ctypes on an n x 1 complex right-hand side. */
typedef BfMat *(*RefGmresFn)(BfMat const *, BfMat const *, BfMat *, BfReal, BfSize, BfSize *, BfMat const *);
double refCallGMRESPrecond(RefGmresFn fn, BfMat const *A, BfMat const *M, const double *b, int64_t n, double tol, int64_t maxNumIter, double *x, int64_t *numIter);
double refCallGMRES(RefGmresFn fn, BfMat const *A, const double *b, int64_t n, double tol, int64_t maxNumIter, double *x, int64_t *numIter) {
  return refCallGMRESPrecond(fn, A, NULL, b, n, tol, maxNumIter, x, numIter);
}

/* The block Jacobi preconditioner of multiple_scattering_context.c:544-600: LU factorizations of the
 * diagonal blocks of the dense second-kind matrix K diag(w) + I/2 (tree order), collected as a
 * MatBlockDiag of bfLuGetMatView objects whose Solve is the LU solve.  `weight` as refHelm2MakeSecondKind. */
BfMat *refHelm2MakeBlockJacobi(RefHelm2 *h, int64_t numBlocks, double weight) {
  BfTree *tree = bfQuadtreeToTree(&h->quadtree);
  BfPerm *revPerm = bfPermGetReversePerm(bfTreeGetPermConst(tree));
  BfMat *ptsPerm = bfMatCopy(bfMatDenseRealToMat(h->pointsMat));
  bfMatPermuteRows(ptsPerm, revPerm);
  BfPoints2 const *pp = bfPoints2ConstViewFromMat(ptsPerm);
  BfMat *K = bfHelm2GetKernelMatrix(&h->helm, pp, pp, NULL, NULL);
  BfSize n = bfMatGetNumRows(K);
  BfMatDenseComplex *Kd = bfMatToMatDenseComplex(K);
  for (BfSize i = 0; i < n; ++i) {
    for (BfSize j = 0; j < n; ++j) Kd->data[i * Kd->rowStride + j] *= weight;
    Kd->data[i * Kd->rowStride + i] += 0.5;
  }
  BfPtrArray *blocks = bfPtrArrayNewWithDefaultCapacity();
  for (int64_t b = 0; b < numBlocks; ++b) {
    BfSize i0 = n * (BfSize)b / (BfSize)numBlocks, i1 = n * (BfSize)(b + 1) / (BfSize)numBlocks;
    BfMat *KBlock = bfMatGetBlockView(K, i0, i1, i0, i1);
    BfLu *lu = bfMatGetLu(KBlock);
    bfPtrArrayAppend(blocks, bfLuGetMatView(lu));
    bfMatDelete(&KBlock);
  }
  BfMat *M = bfMatBlockDiagToMat(bfMatBlockDiagNewFromBlocks(blocks, BF_POLICY_STEAL));
  bfMatDelete(&K);
  bfMatDelete(&ptsPerm);
  bfPermDeinitAndDealloc(&revPerm);
  return M; /* leaked by the tests on purpose: the LU objects have no owner to free them through */
}

double refCallGMRESPrecond(RefGmresFn fn, BfMat const *A, BfMat const *M, const double *b, int64_t n, double tol, int64_t maxNumIter, double *x, int64_t *numIter) {
  BfMatDenseComplex *B = bfMatDenseComplexNewViewFromPtr((BfSize)n, 1, (BfComplex *)b);
  BfSize it = 0;
  double t0 = refNow();
  BfMat *X = fn(A, bfMatDenseComplexToMat(B), NULL, tol, (BfSize)maxNumIter, &it, M);
  double t1 = refNow();
  *numIter = X ? (int64_t)it : -1;
  if (X) {
    BfMatDenseComplex *Xd = bfMatToMatDenseComplex(X);
    for (int64_t i = 0; i < n; ++i) memcpy(x + 2 * i, Xd->data + i * Xd->rowStride, sizeof(BfComplex));
    bfMatDelete(&X);
  }
  BfMat *Bm = bfMatDenseComplexToMat(B);
  bfMatDelete(&Bm);
  return t1 - t0;
}

/* ---- the other public entry points of the apply path, for the attach-layer tests ------------------ */

/* Y = bfMatRmul(A, X) = X A through the public vtable (src/mat.c:193; src/mat_product.c:282-310):
 * X is p x m row-major complex (m = rows of A), Y is p x n.  Returns seconds inside bfMatRmul. */
double refMatRmulComplex(BfMat const *A, const double *x, int64_t p, int64_t m, double *y) {
  BfMatDenseComplex *X = bfMatDenseComplexNewViewFromPtr((BfSize)p, (BfSize)m, (BfComplex *)x);
  double t0 = refNow();
  BfMat *Y = bfMatRmul(A, bfMatDenseComplexToMat(X));
  double t1 = refNow();
  BfMatDenseComplex *Yd = bfMatToMatDenseComplex(Y);
  BfSize n = bfMatGetNumCols(Y);
  for (BfSize i = 0; i < (BfSize)p; ++i)
    memcpy(y + 2 * i * n, Yd->data + i * Yd->rowStride, sizeof(BfComplex) * (size_t)n);
  bfMatDelete(&Y);
  BfMat *Xm = bfMatDenseComplexToMat(X);
  bfMatDelete(&Xm);
  return t1 - t0;
}

/* bfMatMulVec / bfMatRmulVec with a STRIDED BfVecReal view (bfVecRealInitView, include/bf/vec_real.h:48):
 * element i of the operand is x[i * stride]. */
double refMatMulVecRealStrided(BfMat const *A, const double *x, int64_t n, int64_t stride, double *y, int transposed) {
  BfVecReal *xv = bfVecRealNew();
  bfVecRealInitView(xv, (BfSize)n, (BfSize)stride, (BfReal *)x);
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

/* bfMatMulVec / bfMatRmulVec with a (strided) BfVecComplex view.  The reference's block types refuse this
 * operand type (src/mat_block_diag.c:418-428): only call it on an ATTACHED object. */
double refMatMulVecComplexStrided(BfMat const *A, const double *x, int64_t n, int64_t stride, double *y, int transposed) {
  BfVecComplex *xv = bfVecComplexNew();
  bfVecComplexInitView(xv, (BfSize)n, (BfSize)stride, (BfComplex *)x);
  double t0 = refNow();
  BfVec *yv = transposed ? bfMatRmulVec(A, bfVecComplexToVec(xv)) : bfMatMulVec(A, bfVecComplexToVec(xv));
  double t1 = refNow();
  if (yv == NULL) return -1.0;
  BfVecComplex *yc = bfVecToVecComplex(yv);
  for (BfSize i = 0; i < yv->size; ++i) ((BfComplex *)y)[i] = yc->data[i * yc->stride];
  bfVecDelete(&yv);
  BfVec *xvv = bfVecComplexToVec(xv);
  bfVecDelete(&xvv);
  return t1 - t0;
}

/* aliasing slots: bfMatGet(A, BF_POLICY_VIEW) as multiple_scattering_context.c:616 takes it, bfMatCopy, bfMatDelete */
BfMat *refMatGetView(BfMat *A) { return bfMatGetView(A); }
BfMat *refMatCopy(BfMat const *A) { return bfMatCopy(A); }
void refMatDelete(BfMat *A) { bfMatDelete(&A); }
/* the typed view constructor called directly (bypasses the vtable): the view is a struct copy of the owner */
BfMat *refMatBlockDenseGetViewDirect(BfMat *A) { return bfMatBlockDenseGetView((BfMatBlockDense *)A); }
void refHelm2SetMat(RefHelm2 *h, BfMat *A) { h->A = A; }
