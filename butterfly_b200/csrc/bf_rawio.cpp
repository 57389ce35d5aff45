// bf_rawio.cpp — flat little-endian file format for a flattened operator (raw ops + values), so a
// factorization that took the host an hour to build is built once.  The reference has bfMatDump
// writers (src/mat.c:67-73, src/mat_product.c:123-140, src/mat_block_coo.c:260-289) but no reader
// and no Dump for complex dense leaves; this format is ours.
//
//   char     magic[8] = "BFRAWOP1"
//   int32    dtype, reserved
//   int64    numRows, numCols, numBufs, numOps, numValueDoubles
//   int64    bufRows[numBufs]
//   RawOp    ops[numOps]        (72 bytes each; `val` = offset in DOUBLES into values, or -1)
//   double   values[numValueDoubles]   (each leaf re-packed with ld == stored column count)
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "bf_cuda.h"

namespace {
struct Header {
  char magic[8];
  int32_t dtype, reserved;
  int64_t numRows, numCols, numBufs, numOps, numValueDoubles;
};
static_assert(sizeof(BfCudaRawOp) == 72, "raw op ABI");

void storedShape(const BfCudaRawOp &o, bool cplx, int64_t *sr, int64_t *sc) {
  const uint32_t mask = cplx ? (BF_CUDA_LEAF_TRANS | BF_CUDA_LEAF_CONJ) : BF_CUDA_LEAF_TRANS;
  const bool st = (o.flags & mask) != 0;
  *sr = st ? o.k : o.m;
  *sc = st ? o.m : o.k;
}
int64_t leafDoubles(const BfCudaRawOp &o, bool cplx) {
  if (!o.val) return 0;
  if (o.kind == BF_CUDA_LEAF_DENSE) return o.m * o.k * (cplx ? 2 : 1);
  if (o.kind == BF_CUDA_LEAF_DIAG) return o.m < o.k ? o.m : o.k;
  return 0;
}
} // namespace

extern "C" {

BfCudaStatus bfCudaRawOpsSave(const BfCudaPlanDesc *d, const char *path) {
  if (!d || !path) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  const bool cplx = d->dtype == BF_CUDA_COMPLEX_F64;
  // A value-less dense / diagonal leaf (an operator whose values were built on the GPU) would reload as
  // a silent all-zero block: refuse.  Such operators are persisted as compiled plans (bfCudaPlanSave).
  for (int64_t i = 0; i < d->numOps; ++i) {
    const BfCudaRawOp &o = d->ops[i];
    if ((o.kind == BF_CUDA_LEAF_DENSE || o.kind == BF_CUDA_LEAF_DIAG) && o.m > 0 && o.k > 0 && !o.val)
      return BF_CUDA_ERROR_INVALID_ARGUMENT;
  }
  FILE *fp = std::fopen(path, "wb");
  if (!fp) return BF_CUDA_ERROR_IO;
  Header h{};
  std::memcpy(h.magic, "BFRAWOP1", 8);
  h.dtype = d->dtype;
  h.numRows = d->numRows; h.numCols = d->numCols; h.numBufs = d->numBufs; h.numOps = d->numOps;
  std::vector<BfCudaRawOp> ops(d->ops, d->ops + d->numOps);
  int64_t off = 0;
  for (auto &o : ops) {
    const int64_t n = leafDoubles(o, cplx);
    if (n == 0) { o.val = (const void *)(intptr_t)-1; continue; }
    o.val = (const void *)(intptr_t)off;
    if (o.kind == BF_CUDA_LEAF_DENSE) { int64_t sr, sc; storedShape(o, cplx, &sr, &sc); o.ld = sc; }
    off += n;
  }
  h.numValueDoubles = off;
  bool ok = std::fwrite(&h, sizeof h, 1, fp) == 1;
  ok = ok && (d->numBufs == 0 || std::fwrite(d->bufRows, 8, (size_t)d->numBufs, fp) == (size_t)d->numBufs);
  ok = ok && (ops.empty() || std::fwrite(ops.data(), sizeof(BfCudaRawOp), ops.size(), fp) == ops.size());
  const int e = cplx ? 2 : 1;
  for (int64_t i = 0; ok && i < d->numOps; ++i) {
    const BfCudaRawOp &o = d->ops[i];
    if (leafDoubles(o, cplx) == 0) continue;
    const double *src = static_cast<const double *>(o.val);
    if (o.kind == BF_CUDA_LEAF_DIAG) {
      const size_t n = (size_t)(o.m < o.k ? o.m : o.k);
      ok = std::fwrite(src, 8, n, fp) == n;
    } else {
      int64_t sr, sc;
      storedShape(o, cplx, &sr, &sc);
      for (int64_t r = 0; ok && r < sr; ++r)
        ok = std::fwrite(src + r * o.ld * e, 8, (size_t)(sc * e), fp) == (size_t)(sc * e);
    }
  }
  ok = (std::fclose(fp) == 0) && ok;
  return ok ? BF_CUDA_OK : BF_CUDA_ERROR_IO;
}

BfCudaStatus bfCudaRawOpsLoad(const char *path, BfCudaPlanDesc **out) {
  if (!path || !out) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  *out = nullptr;
  FILE *fp = std::fopen(path, "rb");
  if (!fp) return BF_CUDA_ERROR_IO;
  Header h;
  if (std::fread(&h, sizeof h, 1, fp) != 1 || std::memcmp(h.magic, "BFRAWOP1", 8) != 0 || h.numBufs < 2 ||
      h.numOps < 0 || h.numValueDoubles < 0 || (h.dtype != BF_CUDA_REAL_F64 && h.dtype != BF_CUDA_COMPLEX_F64)) {
    std::fclose(fp);
    return BF_CUDA_ERROR_IO;
  }
  // every count is bounded by what the file can hold (a corrupt header must not size a malloc)
  {
    long cur = std::ftell(fp);
    std::fseek(fp, 0, SEEK_END);
    const long end = std::ftell(fp);
    std::fseek(fp, cur, SEEK_SET);
    const uint64_t avail = end > cur ? (uint64_t)(end - cur) : 0;
    const uint64_t nb = (uint64_t)h.numBufs, no = (uint64_t)h.numOps, nv = (uint64_t)h.numValueDoubles;
    if (nb > avail / 8 || no > avail / sizeof(BfCudaRawOp) || nv > avail / 8 ||
        nb * 8 + no * sizeof(BfCudaRawOp) + nv * 8 != avail) {
      std::fclose(fp);
      return BF_CUDA_ERROR_IO;
    }
  }
  const size_t szDesc = (sizeof(BfCudaPlanDesc) + 15) & ~size_t(15);
  const size_t szBufs = ((size_t)h.numBufs * 8 + 15) & ~size_t(15);
  const size_t szOps = ((size_t)h.numOps * sizeof(BfCudaRawOp) + 15) & ~size_t(15);
  const size_t szVals = (size_t)h.numValueDoubles * 8;
  char *arena = static_cast<char *>(std::malloc(szDesc + szBufs + szOps + szVals + 16));
  if (!arena) { std::fclose(fp); return BF_CUDA_ERROR_OUT_OF_MEMORY; }
  BfCudaPlanDesc *d = reinterpret_cast<BfCudaPlanDesc *>(arena);
  int64_t *bufRows = reinterpret_cast<int64_t *>(arena + szDesc);
  BfCudaRawOp *ops = reinterpret_cast<BfCudaRawOp *>(arena + szDesc + szBufs);
  double *vals = reinterpret_cast<double *>(arena + szDesc + szBufs + szOps);
  bool ok = std::fread(bufRows, 8, (size_t)h.numBufs, fp) == (size_t)h.numBufs;
  ok = ok && (h.numOps == 0 || std::fread(ops, sizeof(BfCudaRawOp), (size_t)h.numOps, fp) == (size_t)h.numOps);
  ok = ok && (szVals == 0 || std::fread(vals, 1, szVals, fp) == szVals);
  std::fclose(fp);
  const bool cplx = h.dtype == BF_CUDA_COMPLEX_F64;
  for (int64_t i = 0; ok && i < h.numBufs; ++i) ok = bufRows[i] >= 0;
  for (int64_t i = 0; ok && i < h.numOps; ++i) {
    BfCudaRawOp &o = ops[i];
    const intptr_t off = (intptr_t)o.val;
    o.val = nullptr;
    // shape and placement must be sane before the leaf's extent can be computed at all
    if (o.m < 0 || o.k < 0 || o.m > (int64_t(1) << 40) || o.k > (int64_t(1) << 40) || o.inBuf < 0 || o.inBuf >= h.numBufs ||
        o.outBuf < 0 || o.outBuf >= h.numBufs || o.inOff < 0 || o.outOff < 0 ||
        (o.kind != BF_CUDA_LEAF_DENSE && o.kind != BF_CUDA_LEAF_IDENTITY && o.kind != BF_CUDA_LEAF_DIAG)) { ok = false; break; }
    if (off == -1) continue;
    o.val = vals; // non-null so that leafDoubles counts it
    int64_t sr, sc;
    storedShape(o, cplx, &sr, &sc);
    const int64_t need = leafDoubles(o, cplx);
    if (off < 0 || need > h.numValueDoubles || off > h.numValueDoubles - need ||
        (o.kind == BF_CUDA_LEAF_DENSE && o.ld != sc)) { ok = false; break; }
    o.val = vals + off;
  }
  if (!ok) { std::free(arena); return BF_CUDA_ERROR_IO; }
  std::memset(d, 0, sizeof *d);
  d->abiVersion = BF_CUDA_ABI_VERSION;
  d->dtype = h.dtype;
  d->numRows = h.numRows; d->numCols = h.numCols;
  d->numBufs = h.numBufs; d->bufRows = bufRows;
  d->numOps = h.numOps; d->ops = ops;
  d->device = -1;
  *out = d;
  return BF_CUDA_OK;
}

void bfCudaRawOpsFree(BfCudaPlanDesc *d) { std::free(d); }

} // extern "C"
