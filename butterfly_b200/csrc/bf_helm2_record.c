/* bf_helm2_record.c — records the arguments of the reference's two Helmholtz value builders so that
 * the values can be computed on the GPU (bfCudaPlanBuildHelm2).
 *
 * Every leaf of a fac_helm2 operator is the result of ONE call of
 *   bfHelm2GetKernelMatrix       (src/helm2.c:282-319), or
 *   bfHelm2GetReexpansionMatrix  (src/helm2.c:321-365),
 * made from src/fac_helm2.c:117,364,463,758.  This file defines both symbols.  Built as
 * libbfcuda_helm2.so and placed ahead of libbutterfly in the global symbol scope (or compiled into
 * libbutterfly with the two originals renamed), it sees those calls:
 *   - outside a recording (the default) a call is forwarded to the reference's own definition;
 *   - between bfCudaHelm2RecordBegin and bfCudaHelm2RecordEnd the arguments are copied into a recipe
 *     and the caller gets a value-less BfMatDenseComplex VIEW of the right shape whose `data` points
 *     into a reserved, inaccessible address range (one distinct address per leaf, never dereferenced:
 *     fac_helm2 only stores the leaves into block matrices).
 * bfCudaHelm2Resolve then maps the flattened operator's leaves back to their recipes through those
 * addresses, and bfCudaMatAttachHelm2 compiles the plan and has the GPU fill its value pool.
 *
 * Host C against the reference's headers, like bf_flatten.c; no values are computed here.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <link.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>

#include <bf/helm2.h>
#include <bf/mat_dense_complex.h>
#include <bf/points.h>
#include <bf/vectors.h>

#include "bf_cuda_attach.h"

#define ARENA_BYTES ((size_t)1 << 42) /* 4 TiB of address space, PROT_NONE: 2.7e11 complex entries */

struct BfCudaHelm2Recording {
  unsigned char *arena;   /* reservation the leaves' data pointers point into */
  size_t arenaUsed;       /* bytes handed out */
  double *pairs;          /* pool of (x, y) pairs: points and normals, deduplicated */
  int64_t numPairs, capPairs;
  BfCudaHelm2Recipe *recipes;
  size_t *recipeArena;    /* arena offset of every recipe's element (0, 0); ascending */
  int64_t numRecipes, capRecipes;
  /* open-addressing table of the point sets already in the pool */
  struct Slot { uint64_t hash; int64_t first, count; } *slots;
  size_t numSlots, usedSlots;
  int64_t *opRecipe;      /* filled by bfCudaHelm2Resolve */
  double truncationScale; /* 0 = the reference's truncation rule */
  int failed;
};

static struct BfCudaHelm2Recording *g_rec = NULL; /* the reference is single-threaded (src/error.c:5-7) */

typedef BfMat *(*KernelFn)(BfHelm2 const *, BfPoints2 const *, BfPoints2 const *, BfVectors2 const *, BfVectors2 const *);
typedef BfMat *(*ReexpFn)(BfHelm2 const *, BfPoints2 const *, BfPoints2 const *, BfVectors2 const *, BfVectors2 const *,
                          BfPoints2 const *);

static uint64_t hashBytes(const void *p, size_t n) {
  const unsigned char *b = p;
  uint64_t h = 1469598103934665603ull;
  for (size_t i = 0; i < n; ++i) h = (h ^ b[i]) * 1099511628211ull;
  return h ? h : 1;
}

static void growSlots(struct BfCudaHelm2Recording *r) {
  size_t n = r->numSlots ? 2 * r->numSlots : 1024;
  struct Slot *s = calloc(n, sizeof *s);
  if (!s) { r->failed = 1; return; }
  for (size_t i = 0; i < r->numSlots; ++i) {
    if (!r->slots[i].hash) continue;
    size_t j = r->slots[i].hash & (n - 1);
    while (s[j].hash) j = (j + 1) & (n - 1);
    s[j] = r->slots[i];
  }
  free(r->slots);
  r->slots = s;
  r->numSlots = n;
}

/* Index of the first pair of a copy of `count` pairs starting at `xy` (shared with identical earlier sets). */
static int64_t internPairs(struct BfCudaHelm2Recording *r, const double *xy, int64_t count) {
  if (count == 0) return 0;
  if (2 * (r->usedSlots + 1) > r->numSlots) growSlots(r);
  if (r->failed) return 0;
  const size_t bytes = (size_t)count * 2 * sizeof(double);
  const uint64_t h = hashBytes(xy, bytes);
  size_t j = h & (r->numSlots - 1);
  while (r->slots[j].hash) {
    if (r->slots[j].hash == h && r->slots[j].count == count &&
        memcmp(r->pairs + 2 * r->slots[j].first, xy, bytes) == 0)
      return r->slots[j].first;
    j = (j + 1) & (r->numSlots - 1);
  }
  if (r->numPairs + count > r->capPairs) {
    int64_t cap = r->capPairs ? r->capPairs : 4096;
    while (cap < r->numPairs + count) cap *= 2;
    double *p = realloc(r->pairs, (size_t)cap * 2 * sizeof(double));
    if (!p) { r->failed = 1; return 0; }
    r->pairs = p;
    r->capPairs = cap;
  }
  const int64_t first = r->numPairs;
  memcpy(r->pairs + 2 * first, xy, bytes);
  r->numPairs += count;
  r->slots[j].hash = h; r->slots[j].first = first; r->slots[j].count = count;
  ++r->usedSlots;
  return first;
}

static BfCudaHelm2Recipe *newRecipe(struct BfCudaHelm2Recording *r, BfHelm2 const *helm, int kind, int64_t rows,
                                    int64_t cols, BfComplex **data) {
  if (r->numRecipes == r->capRecipes) {
    int64_t cap = r->capRecipes ? 2 * r->capRecipes : 1024;
    BfCudaHelm2Recipe *p = realloc(r->recipes, (size_t)cap * sizeof *p);
    size_t *q = p ? realloc(r->recipeArena, (size_t)cap * sizeof *q) : NULL;
    if (p) r->recipes = p;
    if (q) r->recipeArena = q;
    if (!p || !q) { r->failed = 1; return NULL; }
    r->capRecipes = cap;
  }
  /* every leaf gets at least 16 bytes of the reservation so that addresses stay distinct */
  size_t bytes = (size_t)rows * (size_t)cols * sizeof(BfComplex);
  if (bytes == 0) bytes = sizeof(BfComplex);
  if (r->arenaUsed + bytes > ARENA_BYTES) { r->failed = 1; return NULL; }
  BfCudaHelm2Recipe *c = &r->recipes[r->numRecipes];
  memset(c, 0, sizeof *c);
  c->kind = kind;
  c->layerPot = (int32_t)helm->layerPot;
  c->k = helm->k;
  c->alpha[0] = creal(helm->alpha); c->alpha[1] = cimag(helm->alpha);
  c->beta[0] = creal(helm->beta); c->beta[1] = cimag(helm->beta);
  c->rows = rows; c->cols = cols;
  c->srcNormals = c->tgtNormals = c->equivNormals = -1;
  r->recipeArena[r->numRecipes] = r->arenaUsed;
  *data = (BfComplex *)(r->arena + r->arenaUsed);
  r->arenaUsed += bytes;
  ++r->numRecipes;
  return c;
}

/* The reference's own definition of `name`: the next one after ours in the search order.  For a library
 * on the program's link line (`-lbfcuda_helm2 -lbutterfly`) that is dlsym(RTLD_NEXT).  A library brought
 * in by dlopen (the Python tools) has only its own dependencies in that order — and this one depends on no
 * libbutterfly on purpose — so as a second step every loaded object is asked for a definition that is not
 * ours. */
struct FindNext { const char *name; void *self, *found; };
static int findNextCb(struct dl_phdr_info *info, size_t size, void *data) {
  (void)size;
  struct FindNext *f = data;
  if (f->found || !info->dlpi_name || !info->dlpi_name[0]) return 0;
  void *h = dlopen(info->dlpi_name, RTLD_LAZY | RTLD_NOLOAD);
  if (!h) return 0;
  void *sym = dlsym(h, f->name);
  dlclose(h);
  if (sym && sym != f->self) f->found = sym;
  return 0;
}
static void *forwardTarget(const char *name, void *self) {
  void *sym = dlsym(RTLD_NEXT, name);
  if (sym && sym != self) return sym;
  struct FindNext f = {name, self, NULL};
  dl_iterate_phdr(findNextCb, &f);
  return f.found;
}

BfMat *bfHelm2GetKernelMatrix(BfHelm2 const *helm, BfPoints2 const *Xsrc, BfPoints2 const *Xtgt,
                              BfVectors2 const *Nsrc, BfVectors2 const *Ntgt) {
  struct BfCudaHelm2Recording *r = g_rec;
  if (!r) {
    static KernelFn next;
    if (!next) next = (KernelFn)forwardTarget("bfHelm2GetKernelMatrix", (void *)bfHelm2GetKernelMatrix);
    if (!next) { bfSetError(BF_ERROR_RUNTIME_ERROR); return NULL; } /* no libbutterfly loaded behind us */
    return next(helm, Xsrc, Xtgt, Nsrc, Ntgt);
  }
  /* the argument checks of src/helm2.c:291-295 */
  if ((Nsrc && Nsrc->size != Xsrc->size) || (Ntgt && Ntgt->size != Xtgt->size)) {
    bfSetError(BF_ERROR_INVALID_ARGUMENTS);
    return NULL;
  }
  BfComplex *data = NULL;
  BfCudaHelm2Recipe *c = newRecipe(r, helm, BF_CUDA_HELM2_KERNEL, (int64_t)Xtgt->size, (int64_t)Xsrc->size, &data);
  if (!c) { bfSetError(BF_ERROR_MEMORY_ERROR); return NULL; }
  c->numSrc = (int64_t)Xsrc->size; c->src = internPairs(r, &Xsrc->data[0][0], c->numSrc);
  c->numTgt = (int64_t)Xtgt->size; c->tgt = internPairs(r, &Xtgt->data[0][0], c->numTgt);
  if (Nsrc) c->srcNormals = internPairs(r, &Nsrc->data[0][0], c->numSrc);
  if (Ntgt) c->tgtNormals = internPairs(r, &Ntgt->data[0][0], c->numTgt);
  /* rows = targets, cols = sources (src/helm2.c:100-101) */
  return bfMatDenseComplexToMat(bfMatDenseComplexNewViewFromPtr(Xtgt->size, Xsrc->size, data));
}

BfMat *bfHelm2GetReexpansionMatrix(BfHelm2 const *helm, BfPoints2 const *srcPtsOrig, BfPoints2 const *srcPtsEquiv,
                                   BfVectors2 const *srcNormalsOrig, BfVectors2 const *srcNormalsEquiv,
                                   BfPoints2 const *tgtPts) {
  struct BfCudaHelm2Recording *r = g_rec;
  if (!r) {
    static ReexpFn next;
    if (!next) next = (ReexpFn)forwardTarget("bfHelm2GetReexpansionMatrix", (void *)bfHelm2GetReexpansionMatrix);
    if (!next) { bfSetError(BF_ERROR_RUNTIME_ERROR); return NULL; }
    return next(helm, srcPtsOrig, srcPtsEquiv, srcNormalsOrig, srcNormalsEquiv, tgtPts);
  }
  /* the argument checks of src/helm2.c:331-342 */
  if ((srcNormalsOrig && srcNormalsOrig->size != srcPtsOrig->size) ||
      (srcNormalsEquiv && srcNormalsEquiv->size != srcPtsEquiv->size)) {
    bfSetError(BF_ERROR_INVALID_ARGUMENTS);
    return NULL;
  }
  if (BF_LAYER_POT_USES_TGT_NORMALS[helm->layerPot]) {
    bfSetError(BF_ERROR_RUNTIME_ERROR);
    return NULL;
  }
  BfComplex *data = NULL;
  /* Z_equiv \ Z_orig: (#equivalent sources) x (#original sources) */
  BfCudaHelm2Recipe *c = newRecipe(r, helm, BF_CUDA_HELM2_REEXPANSION, (int64_t)srcPtsEquiv->size,
                                   (int64_t)srcPtsOrig->size, &data);
  if (!c) { bfSetError(BF_ERROR_MEMORY_ERROR); return NULL; }
  c->numSrc = (int64_t)srcPtsOrig->size; c->src = internPairs(r, &srcPtsOrig->data[0][0], c->numSrc);
  c->numEquiv = (int64_t)srcPtsEquiv->size; c->equiv = internPairs(r, &srcPtsEquiv->data[0][0], c->numEquiv);
  c->numTgt = (int64_t)tgtPts->size; c->tgt = internPairs(r, &tgtPts->data[0][0], c->numTgt);
  if (srcNormalsOrig) c->srcNormals = internPairs(r, &srcNormalsOrig->data[0][0], c->numSrc);
  if (srcNormalsEquiv) c->equivNormals = internPairs(r, &srcNormalsEquiv->data[0][0], c->numEquiv);
  return bfMatDenseComplexToMat(bfMatDenseComplexNewViewFromPtr(srcPtsEquiv->size, srcPtsOrig->size, data));
}

BfCudaStatus bfCudaHelm2RecordBegin(void) {
  if (g_rec) return BF_CUDA_ERROR_INVALID_ARGUMENT; /* recordings do not nest */
  struct BfCudaHelm2Recording *r = calloc(1, sizeof *r);
  if (!r) return BF_CUDA_ERROR_OUT_OF_MEMORY;
  r->arena = mmap(NULL, ARENA_BYTES, PROT_NONE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
  if (r->arena == MAP_FAILED) { free(r); return BF_CUDA_ERROR_OUT_OF_MEMORY; }
  g_rec = r;
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaHelm2RecordEnd(BfCudaHelm2Recording **rec) {
  if (!g_rec || !rec) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  *rec = g_rec;
  g_rec = NULL;
  return (*rec)->failed ? BF_CUDA_ERROR_OUT_OF_MEMORY : BF_CUDA_OK;
}

BfCudaStatus bfCudaHelm2RecordingSetTruncationScale(BfCudaHelm2Recording *r, double scale) {
  if (!r || !(scale >= 0.0)) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  r->truncationScale = scale;
  return BF_CUDA_OK;
}

void bfCudaHelm2RecordingFree(BfCudaHelm2Recording *r) {
  if (!r) return;
  if (r == g_rec) g_rec = NULL;
  munmap(r->arena, ARENA_BYTES);
  free(r->pairs); free(r->recipes); free(r->recipeArena); free(r->slots); free(r->opRecipe);
  free(r);
}

BfCudaStatus bfCudaHelm2Resolve(BfCudaHelm2Recording *r, BfCudaPlanDesc *desc, BfCudaHelm2Build *build) {
  if (!r || !desc || !build) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  free(r->opRecipe);
  r->opRecipe = malloc(sizeof(int64_t) * (size_t)(desc->numOps > 0 ? desc->numOps : 1));
  if (!r->opRecipe) return BF_CUDA_ERROR_OUT_OF_MEMORY;
  BfCudaRawOp *ops = (BfCudaRawOp *)desc->ops; /* the descriptor is ours to edit (bfCudaFlatten allocated it) */
  for (int64_t i = 0; i < desc->numOps; ++i) {
    r->opRecipe[i] = -1;
    const unsigned char *v = ops[i].val;
    if (!v || v < r->arena || v >= r->arena + r->arenaUsed) continue;
    const size_t off = (size_t)(v - r->arena);
    int64_t lo = 0, hi = r->numRecipes; /* the recipe whose arena offset is exactly `off` */
    while (hi - lo > 1) {
      const int64_t mid = (lo + hi) / 2;
      if (r->recipeArena[mid] <= off) lo = mid; else hi = mid;
    }
    const BfCudaHelm2Recipe *c = &r->recipes[lo];
    const int trans = (ops[i].flags & BF_CUDA_LEAF_TRANS) != 0;
    const int64_t storedRows = trans ? ops[i].k : ops[i].m, storedCols = trans ? ops[i].m : ops[i].k;
    /* a leaf must be a whole builder result: fac_helm2 never slices one */
    if (r->recipeArena[lo] != off || ops[i].kind != BF_CUDA_LEAF_DENSE || storedRows != c->rows ||
        storedCols != c->cols || ops[i].ld != c->cols)
      return BF_CUDA_ERROR_NOT_IMPLEMENTED;
    r->opRecipe[i] = lo;
    ops[i].val = NULL; /* "no host values": the pool is filled on the device */
  }
  build->numPairs = r->numPairs;
  build->pairs = r->pairs;
  build->numRecipes = r->numRecipes;
  build->recipes = r->recipes;
  build->numOps = desc->numOps;
  build->opRecipe = r->opRecipe;
  build->truncationScale = r->truncationScale;
  build->opRowOffset = NULL;    /* every leaf is a whole builder result */
  return BF_CUDA_OK;
}
