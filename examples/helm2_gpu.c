/* helm2_gpu.c — the drop-in used from C, the reference's own language: the program of the reference's
 * examples/simple/bf_all_blocks.c (points on a circle -> quadtree -> bfFacHelm2MakeMultilevel -> bfMatMul on a
 * Randn right-hand side -> relative l2 error against the dense kernel matrix) with the GPU behind it.
 * The only lines that are not the reference's API are the bfCuda* ones:
 *
 *   bfCudaHelm2RecordBegin / End   the reference lays out the operator, no values are computed on the host
 *   bfCudaMatAttachHelm2           values computed on the B200, vtable swapped
 *   bfMatMul(A, X)                 unchanged call, runs the sm_100a stage kernels
 *
 * Build (link order matters: libbfcuda_helm2 must precede libbutterfly so that its two builders win):
 *   gcc -std=gnu11 -DBF_DOUBLE -DBF_LINUX -I<butterfly>/include -I<repo>/include examples/helm2_gpu.c \
 *       -L<repo>/butterfly_b200/lib -lbfcuda_helm2 -lbfcuda_attach -lbfcuda -L<butterfly-build> -lbutterfly -lm
 * usage: helm2_gpu [N [k [truncationScale]]]     (defaults 4096, N / 40.96, 0 = the reference's rule)
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>

#include <bf/fac_helm2.h>
#include <bf/helm2.h>
#include <bf/mat_dense_complex.h>
#include <bf/mat_dense_real.h>
#include <bf/points.h>
#include <bf/quadtree.h>
#include <bf/rand.h>

#include "bf_cuda_attach.h"

static double now(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int main(int argc, char **argv) {
  BfSize n = argc > 1 ? (BfSize)atoll(argv[1]) : 4096;
  BfReal k = argc > 2 ? atof(argv[2]) : (BfReal)n / 40.96;
  double scale = argc > 3 ? atof(argv[3]) : 0.0;

  /* examples/simple/make_circle_test_data.py:12-15 */
  BfReal *xy = malloc(sizeof(BfReal) * 2 * n);
  for (BfSize i = 0; i < n; ++i) {
    xy[2 * i] = cos(2 * M_PI * (BfReal)i / (BfReal)n);
    xy[2 * i + 1] = sin(2 * M_PI * (BfReal)i / (BfReal)n);
  }
  BfMatDenseReal *pointsMat = bfMatDenseRealNewViewFromPtr(n, 2, xy, 2, 1);
  BfPoints2 const *points = bfPoints2ConstViewFromMatDenseReal(pointsMat);
  BfQuadtree tree;
  bfQuadtreeInit(&tree, points, NULL);
  BfHelm2 helm = {.k = k, .layerPot = BF_LAYER_POTENTIAL_SINGLE};

  double t0 = now();
  if (bfCudaHelm2RecordBegin() != BF_CUDA_OK) return 2;
  BfMat *A = bfFacHelm2MakeMultilevel(&helm, &tree, &tree); /* structure only */
  BfCudaHelm2Recording *rec = NULL;
  if (bfCudaHelm2RecordEnd(&rec) != BF_CUDA_OK) return 2;
  double t1 = now();
  bfCudaHelm2RecordingSetTruncationScale(rec, scale);
  BfCudaHelm2Stats stats;
  BfCudaStatus st = bfCudaMatAttachHelm2(A, rec, NULL, &stats);
  if (st != BF_CUDA_OK) {
    fprintf(stderr, "bfCudaMatAttachHelm2: %s (%s)\n", bfCudaStatusString(st), bfCudaLastErrorMessage());
    return 1;
  }
  bfCudaHelm2RecordingFree(rec);
  double t2 = now();
  printf("laid out in %.3f s, values built on the GPU and attached in %.3f s (%lld kernel + %lld re-expansion leaves)\n",
         t1 - t0, t2 - t1, (long long)stats.kernelRecipes, (long long)stats.reexpRecipes);

  bfSeed(0);
  BfMat *X = bfMatDenseComplexToMat(bfMatDenseComplexNewRandn(n, 1));
  double t3 = now();
  BfMat *Y = bfMatMul(A, X); /* the reference's call; the GPU's work */
  printf("multiplied with HODBF matrix [%0.4fs]\n", now() - t3);

  /* bf_all_blocks.c:104-153 checks against the whole dense matrix; 64 of its rows do here (tree order) */
  BfPerm const *perm = bfTreeGetPermConst(bfQuadtreeToTree(&tree));
  BfComplex const *x = bfMatToMatDenseComplex(X)->data, *y = bfMatToMatDenseComplex(Y)->data;
  double num = 0, den = 0;
  for (BfSize s = 0; s < 64 && s < n; ++s) {
    BfSize i = (s * 2654435761u) % n;
    BfReal const *pi = &xy[2 * perm->index[i]];
    BfComplex want = 0;
    for (BfSize j = 0; j < n; ++j) {
      BfReal const *pj = &xy[2 * perm->index[j]];
      BfReal r = hypot(pj[0] - pi[0], pj[1] - pi[1]);
      if (r > 0) want += (I * (j0(k * r) + I * y0(k * r)) / 4) * x[j]; /* src/helm2.c:38-41 */
    }
    num += pow(cabs(y[i] - want), 2);
    den += pow(cabs(want), 2);
  }
  printf("relative l2 error: %g\n", sqrt(num / den));
  bfMatDelete(&Y);
  bfMatDelete(&X);
  return sqrt(num / den) < 1e-9 ? 0 : 3;
}
