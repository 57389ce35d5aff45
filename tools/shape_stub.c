/* shape_stub.c — WORKLOAD GENERATION ONLY (bench.py, tools/): lets the unmodified reference's own
 * tree + fac_helm2 logic (src/fac_helm2.c:653-1002, src/quadtree.c) lay out the block STRUCTURE
 * of a Helmholtz butterfly at sizes where computing the factor VALUES would take the reference an
 * hour and ~75 GB of host memory.
 *
 * Loaded with RTLD_GLOBAL *before* libbutterfly_ref.so, these two definitions interpose the
 * reference's kernel-matrix builders (src/helm2.c:282-365): same shapes, no values (data == NULL,
 * VIEW property so nothing is freed).  The flattener then emits value-less raw ops and the pool is
 * filled on the device by bfCudaPlanFillSynthetic.  Nothing numeric of the reference is replaced
 * in the oracle or in any parity test: those always load the real library alone.
 */
#include <bf/helm2.h>
#include <bf/mat_dense_complex.h>
#include <bf/points.h>

BfMat *bfHelm2GetKernelMatrix(BfHelm2 const *helm, BfPoints2 const *Xsrc, BfPoints2 const *Xtgt,
                              BfVectors2 const *Nsrc, BfVectors2 const *Ntgt) {
  (void)helm; (void)Nsrc; (void)Ntgt;
  /* rows = targets, cols = sources (src/helm2.c:282-319) */
  return bfMatDenseComplexToMat(bfMatDenseComplexNewViewFromPtr(Xtgt->size, Xsrc->size, NULL));
}

BfMat *bfHelm2GetReexpansionMatrix(BfHelm2 const *helm, BfPoints2 const *srcPtsOrig,
                                   BfPoints2 const *srcPtsEquiv, BfVectors2 const *srcNormalsOrig,
                                   BfVectors2 const *srcNormalsEquiv, BfPoints2 const *tgtPts) {
  (void)helm; (void)srcNormalsOrig; (void)srcNormalsEquiv; (void)tgtPts;
  /* Z_equiv \ Z_orig: (#equivalent sources) x (#original sources) (src/helm2.c:321-365) */
  return bfMatDenseComplexToMat(bfMatDenseComplexNewViewFromPtr(srcPtsEquiv->size, srcPtsOrig->size, NULL));
}
