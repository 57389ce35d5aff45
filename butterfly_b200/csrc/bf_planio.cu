// bf_planio.cu — persistence of a COMPILED plan: stage tables, single-launch tables and the packed factor pool as it
// sits in HBM.  The reference has writers without a reader (bfMatDump: src/mat.c:67-73,
// src/mat_product.c:123-140, src/mat_block_coo.c:260-289) and no Dump at all for complex dense leaves, so an
// operator that took its host an hour to factorize is rebuilt by every run.  bfCudaRawOpsSave / Load (bf_rawio.cpp)
// store the flattened operator with HOST values; this file stores what the kernels consume, which also covers
// operators whose values only ever existed on the GPU (bfCudaPlanBuildHelm2) and the per-rank plans of a block-row
// sharded operator.
//
//   "BFPLAN02" | header | permutations | per direction: scalars, per stage: tables | single-launch tables | pool bytes
//
// Little-endian, no padding tricks: every array is (u64 count, raw elements).  The pool moves through two pinned
// 64 MiB buffers so that disk and PCIe overlap.  A loaded plan applies exactly like the one that was saved (same
// tables, same bytes: bit-identical results); it holds no per-leaf host data, so its values cannot be re-built
// (bfCudaPlanFillSynthetic / bfCudaPlanBuildHelm2 return NOT_IMPLEMENTED).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "bf_internal.hpp"

using namespace bfc;

namespace {
constexpr char kMagic[8] = {'B', 'F', 'P', 'L', 'A', 'N', '0', '2'};
constexpr size_t kIoChunk = size_t(64) << 20;

struct Writer {
  FILE *fp;
  bool ok = true;
  void raw(const void *p, size_t n) { if (ok && n && std::fwrite(p, 1, n, fp) != n) ok = false; }
  template <class T> void pod(const T &v) { raw(&v, sizeof v); }
  template <class T> void vec(const std::vector<T> &v) { pod((uint64_t)v.size()); raw(v.data(), v.size() * sizeof(T)); }
};
struct Reader {
  FILE *fp;
  uint64_t left; // bytes of the file not consumed yet: no count read from the file may exceed it
  bool ok = true;
  void raw(void *p, size_t n) {
    if (!ok) return;
    if (n > left || (n && std::fread(p, 1, n, fp) != n)) { ok = false; return; }
    left -= n;
  }
  template <class T> void pod(T &v) { raw(&v, sizeof v); }
  template <class T> void vec(std::vector<T> &v) {
    uint64_t n = 0;
    pod(n);
    if (!ok || n > left / sizeof(T)) { ok = false; return; }
    v.resize((size_t)n);
    raw(v.data(), (size_t)n * sizeof(T));
  }
};

struct DirScalars { // fixed-size part of a Direction
  int32_t op, dtype, numQueues, shardWorld, shardRank, numStages;
  int64_t rowsIn, rowsOut, inBufId, outBufId, wsRows, wsArena0, wsArena1, entries, numItems, numTiles, maxTileK, maxItemRows;
  uint64_t pool16;
  int32_t fusedEnabled;
  uint32_t fusedNumCounters;
  int64_t fusedWsRows;
  uint64_t fNumItems, fNumTiles;
};
struct StageScalars {
  uint64_t poolBegin16, poolEnd16;
  int64_t entries, sentRows;
  uint32_t sendMask, recvMask;
};

template <class T> bool fromDevice(std::vector<T> &v, const T *d, size_t n) {
  v.resize(n);
  return n == 0 || cudaMemcpy(v.data(), d, n * sizeof(T), cudaMemcpyDeviceToHost) == cudaSuccess;
}
} // namespace

extern "C" BfCudaStatus bfCudaPlanSave(BfCudaPlan *P, const char *path) {
  if (!P || !path) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  DeviceGuard g(P->device);
  BF_CUDA_TRY(cudaStreamSynchronize(P->stream));
  FILE *fp = std::fopen(path, "wb");
  if (!fp) return fail(BF_CUDA_ERROR_IO, std::string("cannot create ") + path);
  Writer w{fp};
  w.raw(kMagic, 8);
  const int32_t abi = BF_CUDA_ABI_VERSION, dtype = P->dtype;
  w.pod(abi); w.pod(dtype);
  w.pod(P->numRows); w.pod(P->numCols); w.pod(P->numOps);
  // fused caller-side permutations (scatter form; the inverses are rebuilt on load)
  std::vector<int32_t> inPerm, outPerm;
  bool dev = (!P->inPerm || fromDevice(inPerm, P->inPerm, (size_t)P->numCols)) && (!P->outPerm || fromDevice(outPerm, P->outPerm, (size_t)P->numRows));
  w.vec(inPerm); w.vec(outPerm);
  int32_t numDirs = 0;
  for (const DevDirection &dd : P->dir) numDirs += dd.built ? 1 : 0;
  w.pod(numDirs);
  unsigned char *pin[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  for (int i = 0; i < 2 && dev; ++i) dev = cudaHostAlloc(&pin[i], kIoChunk, cudaHostAllocDefault) == cudaSuccess && cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) == cudaSuccess;
  for (const DevDirection &dd : P->dir) {
    if (!dd.built || !dev || !w.ok) continue;
    const Direction &D = dd.host;
    DirScalars s{};
    s.op = D.op; s.dtype = D.dtype; s.numQueues = D.numQueues; s.shardWorld = D.shardWorld; s.shardRank = D.shardRank;
    s.numStages = (int32_t)D.stages.size();
    s.rowsIn = D.rowsIn; s.rowsOut = D.rowsOut; s.inBufId = D.inBufId; s.outBufId = D.outBufId;
    s.wsRows = D.wsRows; s.wsArena0 = D.wsArena[0]; s.wsArena1 = D.wsArena[1];
    s.entries = D.entries; s.numItems = D.numItems; s.numTiles = D.numTiles; s.maxTileK = D.maxTileK; s.maxItemRows = D.maxItemRows;
    s.pool16 = D.pool16;
    s.fusedEnabled = dd.fusedReady ? 1 : 0; s.fusedNumCounters = dd.numCounters; s.fusedWsRows = dd.fusedWsRows;
    s.fNumItems = dd.fNumItems; s.fNumTiles = dd.fNumTiles;
    w.pod(s);
    for (const Stage &S : D.stages) {
      StageScalars ss{S.poolBegin16, S.poolEnd16, S.entries, S.sentRows, S.sendMask, S.recvMask};
      w.pod(ss);
      w.vec(S.items); w.vec(S.tiles); w.vec(S.qItemBegin); w.vec(S.qTileBegin); w.vec(S.qPool16); w.vec(S.itemPeers);
    }
    if (dd.fusedReady) { // the host copies were dropped after the upload: read them back
      std::vector<DevItem> fi; std::vector<ItemDep> fid; std::vector<DevTile> ft; std::vector<TileDep> ftd;
      dev = fromDevice(fi, dd.fItems, dd.fNumItems) && fromDevice(fid, dd.fItemDeps, dd.fNumItems) &&
            fromDevice(ft, dd.fTiles, dd.fNumTiles) && fromDevice(ftd, dd.fTileDeps, dd.fNumTiles);
      w.vec(fi); w.vec(fid); w.vec(ft); w.vec(ftd);
      w.vec(D.fused.qItemBegin); w.vec(D.fused.qTileBegin);
    }
    // the pool, double-buffered: the D2H of chunk c + 1 runs while chunk c is written
    const size_t bytes = (size_t)D.pool16 * 16;
    const size_t nChunks = (bytes + kIoChunk - 1) / kIoChunk;
    auto issue = [&](size_t c) {
      const size_t off = c * kIoChunk, n = std::min(kIoChunk, bytes - off);
      return cudaMemcpyAsync(pin[c & 1], dd.pool + off, n, cudaMemcpyDeviceToHost, P->stream) == cudaSuccess &&
             cudaEventRecord(ev[c & 1], P->stream) == cudaSuccess;
    };
    if (nChunks && dev) dev = issue(0);
    for (size_t c = 0; c < nChunks && dev && w.ok; ++c) {
      if (c + 1 < nChunks) dev = issue(c + 1);
      if (cudaEventSynchronize(ev[c & 1]) != cudaSuccess) { dev = false; break; }
      w.raw(pin[c & 1], std::min(kIoChunk, bytes - c * kIoChunk));
    }
  }
  cudaStreamSynchronize(P->stream);
  for (int i = 0; i < 2; ++i) { if (pin[i]) cudaFreeHost(pin[i]); if (ev[i]) cudaEventDestroy(ev[i]); }
  const bool closed = std::fclose(fp) == 0;
  if (!dev) { cudaGetLastError(); std::remove(path); return fail(BF_CUDA_ERROR_RUNTIME, "device read-back failed while saving the plan"); }
  if (!w.ok || !closed) { std::remove(path); return fail(BF_CUDA_ERROR_IO, std::string("write error on ") + path); }
  return BF_CUDA_OK;
}

namespace {
// Every index a kernel will follow must stay inside what the file itself declares: a truncated or corrupted file
// is refused here, not discovered as an illegal address on the GPU.
bool rowsInside(const Direction &D, uint32_t addr, int64_t rows) {
  const uint32_t sp = addr >> 30;
  const int64_t r = (int64_t)(addr & kRowMask);
  const int64_t lim = sp == kSpaceX ? D.rowsIn : sp == kSpaceY ? D.rowsOut : sp == kSpaceW ? std::max(D.wsRows, D.fused.wsRows) : -1;
  return rows >= 0 && r + rows <= lim;
}
bool validTables(const Direction &D, const std::vector<DevItem> &items, const std::vector<DevTile> &tiles, bool fusedTables) {
  const int maxRows = maxItemRows(D.dtype);
  for (const DevItem &it : items) {
    if ((uint64_t)it.tileBegin + it.numTiles > tiles.size()) return false;
    if (fusedTables && (it.flags & kItemJump)) continue;
    if (it.rows == 0 || it.rows > maxRows || (it.outAddr >> 30) == kSpaceX || !rowsInside(D, it.outAddr, it.rows)) return false;
  }
  const int maxKc = D.dtype == BF_CUDA_COMPLEX_F64 ? kMaxKcComplex : kMaxKcReal;
  for (const DevTile &t : tiles) {
    if (t.kc == kKcJump) { if (!fusedTables || ((((uint64_t)t.rows) << 32) | t.inAddr) > D.pool16) return false; continue; }
    if (t.rows == 0 || t.rows > maxRows) return false;
    const int64_t k = t.kc >= kKcDiag ? t.rows : t.kc;
    if (t.kc < kKcDiag && (t.kc == 0 || t.kc > maxKc)) return false;
    if (!rowsInside(D, t.inAddr, k)) return false;
  }
  return true;
}
} // namespace

extern "C" BfCudaStatus bfCudaPlanLoad(const char *path, int32_t device, BfCudaPlan **out) {
  if (!path || !out) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  *out = nullptr;
  int nDev = 0;
  if (cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0) {
    cudaGetLastError();
    return fail(BF_CUDA_ERROR_NO_DEVICE, "no CUDA device visible; this library has no CPU fallback");
  }
  int dev = device;
  if (dev < 0) BF_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= nDev) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "device ordinal out of range");
  DeviceGuard g(dev);
  if (!g.ok) return fail(BF_CUDA_ERROR_RUNTIME, "cudaSetDevice failed");
  cudaDeviceProp prop;
  BF_CUDA_TRY(cudaGetDeviceProperties(&prop, dev));
  if (prop.major != 10) return fail(BF_CUDA_ERROR_NO_DEVICE, "device is not sm_100-class; kernels are built for sm_100a only");
  BF_CUDA_TRY(configureKernels());
  FILE *fp = std::fopen(path, "rb");
  if (!fp) return fail(BF_CUDA_ERROR_IO, std::string("cannot open ") + path);
  std::fseek(fp, 0, SEEK_END);
  const long fileSize = std::ftell(fp);
  std::fseek(fp, 0, SEEK_SET);
  Reader r{fp, fileSize > 0 ? (uint64_t)fileSize : 0};
  std::unique_ptr<BfCudaPlan> P(new BfCudaPlan());
  unsigned char *pin[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  auto bail = [&](BfCudaStatus st, const std::string &msg) {
    std::fclose(fp);
    for (int i = 0; i < 2; ++i) { if (pin[i]) cudaFreeHost(pin[i]); if (ev[i]) cudaEventDestroy(ev[i]); }
    bfCudaPlanDestroy(P.release());
    return fail(st, msg);
  };
  char magic[8];
  int32_t abi = 0, dtype = 0, numDirs = 0;
  r.raw(magic, 8); r.pod(abi); r.pod(dtype);
  if (!r.ok || std::memcmp(magic, kMagic, 8) != 0) return bail(BF_CUDA_ERROR_IO, "not a compiled-plan file (BFPLAN02)");
  if (abi != BF_CUDA_ABI_VERSION) return bail(BF_CUDA_ERROR_IO, "plan file written by another ABI version");
  if (dtype != BF_CUDA_REAL_F64 && dtype != BF_CUDA_COMPLEX_F64) return bail(BF_CUDA_ERROR_IO, "corrupt plan file (dtype)");
  P->device = dev; P->numSMs = prop.multiProcessorCount; P->cooperativeOk = prop.cooperativeLaunch != 0;
  P->dtype = dtype; P->loadedFromFile = true;
  r.pod(P->numRows); r.pod(P->numCols); r.pod(P->numOps);
  std::vector<int32_t> inPerm, outPerm;
  r.vec(inPerm); r.vec(outPerm);
  r.pod(numDirs);
  if (!r.ok || P->numRows < 0 || P->numCols < 0 || numDirs < 1 || numDirs > 3 || (!inPerm.empty() && (int64_t)inPerm.size() != P->numCols) ||
      (!outPerm.empty() && (int64_t)outPerm.size() != P->numRows))
    return bail(BF_CUDA_ERROR_IO, "corrupt plan file (header)");
  if (cudaStreamCreateWithFlags(&P->stream, cudaStreamNonBlocking) != cudaSuccess) return bail(BF_CUDA_ERROR_RUNTIME, "stream");
  for (int i = 0; i < 2; ++i)
    if (cudaHostAlloc(&pin[i], kIoChunk, cudaHostAllocDefault) != cudaSuccess || cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) != cudaSuccess)
      return bail(BF_CUDA_ERROR_OUT_OF_MEMORY, "pinned staging buffers");
  for (int d = 0; d < numDirs; ++d) {
    DirScalars s{};
    r.pod(s);
    if (!r.ok || s.op < 0 || s.op > 2 || P->dir[s.op].built || s.dtype != dtype || s.numQueues <= 0 || s.numQueues > (1 << 20) ||
        s.numStages < 0 || s.numStages > 4096 || s.rowsIn != (s.op == BF_CUDA_OP_N ? P->numCols : P->numRows) ||
        s.rowsOut != (s.op == BF_CUDA_OP_N ? P->numRows : P->numCols) || s.wsRows < 0 || s.wsRows > (int64_t)kRowMask ||
        s.fusedWsRows < 0 || s.fusedWsRows > (int64_t)kRowMask || s.shardWorld < 0 || s.shardWorld > 16 || s.pool16 > r.left / 16)
      return bail(BF_CUDA_ERROR_IO, "corrupt plan file (direction header)");
    DevDirection &dd = P->dir[s.op];
    Direction &D = dd.host;
    D.op = s.op; D.dtype = s.dtype; D.numQueues = s.numQueues; D.shardWorld = s.shardWorld; D.shardRank = s.shardRank;
    D.rowsIn = s.rowsIn; D.rowsOut = s.rowsOut; D.inBufId = s.inBufId; D.outBufId = s.outBufId;
    D.wsRows = s.wsRows; D.wsArena[0] = s.wsArena0; D.wsArena[1] = s.wsArena1;
    D.entries = s.entries; D.numItems = s.numItems; D.numTiles = s.numTiles; D.maxTileK = s.maxTileK; D.maxItemRows = s.maxItemRows;
    D.pool16 = s.pool16;
    D.fused.wsRows = s.fusedEnabled ? s.fusedWsRows : 0;
    D.stages.assign((size_t)s.numStages, Stage());
    const size_t Q1 = (size_t)s.numQueues + 1;
    for (Stage &S : D.stages) {
      StageScalars ss{};
      r.pod(ss);
      S.poolBegin16 = ss.poolBegin16; S.poolEnd16 = ss.poolEnd16; S.entries = ss.entries; S.sentRows = ss.sentRows;
      S.sendMask = ss.sendMask; S.recvMask = ss.recvMask;
      r.vec(S.items); r.vec(S.tiles); r.vec(S.qItemBegin); r.vec(S.qTileBegin); r.vec(S.qPool16); r.vec(S.itemPeers);
      bool good = r.ok && S.qItemBegin.size() == Q1 && S.qTileBegin.size() == Q1 && S.qPool16.size() == Q1 &&
                  S.poolBegin16 <= S.poolEnd16 && S.poolEnd16 <= D.pool16 &&
                  S.itemPeers.size() == (D.shardWorld > 1 ? S.items.size() : 0) && validTables(D, S.items, S.tiles, false);
      for (size_t q = 0; good && q < Q1; ++q)
        good = S.qItemBegin[q] <= S.items.size() && S.qTileBegin[q] <= S.tiles.size() && S.qPool16[q] <= D.pool16 &&
               (q == 0 || (S.qItemBegin[q] >= S.qItemBegin[q - 1] && S.qTileBegin[q] >= S.qTileBegin[q - 1]));
      if (!good) return bail(BF_CUDA_ERROR_IO, "corrupt plan file (stage tables)");
    }
    if (s.fusedEnabled) {
      FusedTables &F = D.fused;
      r.vec(F.items); r.vec(F.itemDeps); r.vec(F.tiles); r.vec(F.tileDeps); r.vec(F.qItemBegin); r.vec(F.qTileBegin);
      F.numCounters = s.fusedNumCounters;
      bool good = r.ok && F.items.size() == s.fNumItems && F.itemDeps.size() == F.items.size() && F.tiles.size() == s.fNumTiles &&
                  F.tileDeps.size() == F.tiles.size() && F.qItemBegin.size() == Q1 && F.qTileBegin.size() == Q1 && D.shardWorld <= 1 &&
                  validTables(D, F.items, F.tiles, true);
      for (size_t i = 0; good && i < F.itemDeps.size(); ++i) good = F.itemDeps[i].sig == kNoCounter || F.itemDeps[i].sig < F.numCounters;
      for (size_t i = 0; good && i < F.tileDeps.size(); ++i) good = F.tileDeps[i].idx == kNoCounter || F.tileDeps[i].idx < F.numCounters;
      for (size_t q = 0; good && q < Q1; ++q)
        good = F.qItemBegin[q] <= F.items.size() && F.qTileBegin[q] <= F.tiles.size() && (q == 0 || F.qItemBegin[q] >= F.qItemBegin[q - 1]);
      if (!good) return bail(BF_CUDA_ERROR_IO, "corrupt plan file (single-launch tables)");
      F.enabled = true;
    }
    const BfCudaStatus st = uploadDirection(P.get(), dd); // tables to the device, a zeroed pool
    if (st != BF_CUDA_OK) return bail(st, bfCudaLastErrorMessage());
    const size_t bytes = (size_t)D.pool16 * 16;
    for (size_t c = 0, off = 0; off < bytes; ++c, off += kIoChunk) { // disk -> pinned -> HBM, two buffers in flight
      const size_t n = std::min(kIoChunk, bytes - off);
      if (c >= 2 && cudaEventSynchronize(ev[c & 1]) != cudaSuccess) return bail(BF_CUDA_ERROR_RUNTIME, "pool upload");
      r.raw(pin[c & 1], n);
      if (!r.ok) return bail(BF_CUDA_ERROR_IO, "plan file ends inside the factor pool");
      if (cudaMemcpyAsync(dd.pool + off, pin[c & 1], n, cudaMemcpyHostToDevice, P->stream) != cudaSuccess ||
          cudaEventRecord(ev[c & 1], P->stream) != cudaSuccess)
        return bail(BF_CUDA_ERROR_RUNTIME, "pool upload");
    }
    if (cudaStreamSynchronize(P->stream) != cudaSuccess) return bail(BF_CUDA_ERROR_RUNTIME, "pool upload");
  }
  if (!P->dir[0].built || r.left != 0) return bail(BF_CUDA_ERROR_IO, "corrupt plan file (no forward direction, or trailing bytes)");
  std::fclose(fp);
  for (int i = 0; i < 2; ++i) { cudaFreeHost(pin[i]); cudaEventDestroy(ev[i]); }
  if (!inPerm.empty() || !outPerm.empty()) {
    std::vector<int64_t> a(inPerm.begin(), inPerm.end()), b(outPerm.begin(), outPerm.end());
    const BfCudaStatus st = bfCudaPlanSetPerms(P.get(), a.empty() ? nullptr : a.data(), b.empty() ? nullptr : b.data());
    if (st != BF_CUDA_OK) { bfCudaPlanDestroy(P.release()); return st; }
  }
  *out = P.release();
  return BF_CUDA_OK;
}
