// bf_capi.cu — the C ABI of include/bf_cuda.h: plan life cycle, device / host apply, synthetic
// fill, profiling.  No CPU fallback anywhere: every entry point that computes needs a CUDA device.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "bf_internal.hpp"

using namespace bfc;

namespace {
thread_local std::string g_lastError;
} // namespace
namespace bfc {
BfCudaStatus fail(BfCudaStatus st, const std::string &msg) { g_lastError = msg; return st; }
} // namespace bfc

namespace bfc {
BfCudaStatus uploadDirection(BfCudaPlan *P, DevDirection &dd) {
  Direction &D = dd.host;
  const size_t nS = D.stages.size();
  const size_t Q = (size_t)D.numQueues;
  size_t nI = 0, nT = 0;
  dd.itemOff.assign(nS + 1, 0);
  dd.tileOff.assign(nS + 1, 0);
  for (size_t s = 0; s < nS; ++s) {
    dd.itemOff[s] = nI; dd.tileOff[s] = nT;
    nI += D.stages[s].items.size(); nT += D.stages[s].tiles.size();
  }
  dd.itemOff[nS] = nI; dd.tileOff[nS] = nT;
  BF_CUDA_TRY(cudaMalloc(&dd.items, std::max<size_t>(nI, 1) * sizeof(DevItem)));
  BF_CUDA_TRY(cudaMalloc(&dd.tiles, std::max<size_t>(nT, 1) * sizeof(DevTile)));
  BF_CUDA_TRY(cudaMalloc(&dd.qItemBegin, nS * (Q + 1) * sizeof(uint32_t)));
  BF_CUDA_TRY(cudaMalloc(&dd.qTileBegin, nS * (Q + 1) * sizeof(uint32_t)));
  BF_CUDA_TRY(cudaMalloc(&dd.qPool16, nS * (Q + 1) * sizeof(uint64_t)));
  dd.tableBytes = (int64_t)(nI * sizeof(DevItem) + nT * sizeof(DevTile) + nS * (Q + 1) * 16);
  for (size_t s = 0; s < nS; ++s) {
    const Stage &S = D.stages[s];
    if (!S.items.empty())
      BF_CUDA_TRY(cudaMemcpy(dd.items + dd.itemOff[s], S.items.data(), S.items.size() * sizeof(DevItem), cudaMemcpyHostToDevice));
    if (!S.tiles.empty())
      BF_CUDA_TRY(cudaMemcpy(dd.tiles + dd.tileOff[s], S.tiles.data(), S.tiles.size() * sizeof(DevTile), cudaMemcpyHostToDevice));
    BF_CUDA_TRY(cudaMemcpy(dd.qItemBegin + s * (Q + 1), S.qItemBegin.data(), (Q + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice));
    BF_CUDA_TRY(cudaMemcpy(dd.qTileBegin + s * (Q + 1), S.qTileBegin.data(), (Q + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice));
    BF_CUDA_TRY(cudaMemcpy(dd.qPool16 + s * (Q + 1), S.qPool16.data(), (Q + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice));
  }
  const size_t poolBytes = (size_t)D.pool16 * 16 + 256; // slack: nothing reads it, keeps zero-size plans valid
  BF_CUDA_TRY(cudaMalloc(&dd.pool, poolBytes));
  BF_CUDA_TRY(cudaMemsetAsync(dd.pool, 0, poolBytes, P->stream)); // (ops without host values stay zero)
  bool anyVal = false;
  for (const DirOp &d : D.ops) if (d.val) { anyVal = true; break; }
  if (anyVal) {
    // Pack chunk by chunk (whole tiles, <= kPackChunk bytes) into two pinned buffers: the host threads pack chunk
    // c + 1 while the DMA engine moves chunk c.  (A single thread through an unpinned buffer with blocking copies
    // moved 2 GB/s: 1.65 s to attach the 3.3 GB operator of N = 65 536.)
    constexpr size_t kPackChunk = size_t(32) << 20;
    unsigned char *pin[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    cudaStream_t up = P->stream;
    auto cleanup = [&]() { for (int i = 0; i < 2; ++i) { if (pin[i]) cudaFreeHost(pin[i]); if (ev[i]) cudaEventDestroy(ev[i]); } };
    for (int i = 0; i < 2; ++i)
      if (cudaHostAlloc(&pin[i], kPackChunk + kTileBytes, cudaHostAllocDefault) != cudaSuccess || cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) != cudaSuccess) {
        cleanup();
        return fail(BF_CUDA_ERROR_OUT_OF_MEMORY, "pinned staging buffers for the value pool");
      }
    const unsigned nThreads = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    size_t chunkNo = 0;
    cudaError_t ce = cudaSuccess;
    for (size_t s = 0; s < nS && ce == cudaSuccess; ++s) {
      const Stage &S = D.stages[s];
      const size_t nT = S.tiles.size();
      for (size_t t0 = 0; t0 < nT && ce == cudaSuccess;) {
        const uint64_t base16 = S.tilePool16[t0];
        size_t t1 = t0;
        uint64_t end16 = base16;
        while (t1 < nT) { // whole tiles up to the chunk size
          const uint64_t e = S.tilePool16[t1] + tileBytes16(D.dtype, S.tiles[t1].rows, S.tiles[t1].kc);
          if ((e - base16) * 16 > kPackChunk && t1 > t0) break;
          end16 = e;
          ++t1;
        }
        const size_t bytes = (size_t)(end16 - base16) * 16;
        if (bytes) {
          const int b = (int)(chunkNo & 1);
          if (chunkNo >= 2) ce = cudaEventSynchronize(ev[b]); // the copy out of this buffer has finished
          std::memset(pin[b], 0, bytes);
          std::vector<std::thread> th;
          const size_t per = (t1 - t0 + nThreads - 1) / nThreads;
          for (unsigned k = 0; k < nThreads && t0 + k * per < t1; ++k)
            th.emplace_back([&, k]() { packTileRange(D, s, t0 + k * per, std::min(t1, t0 + (k + 1) * per), base16, pin[b]); });
          for (std::thread &x : th) x.join();
          if (ce == cudaSuccess) ce = cudaMemcpyAsync(dd.pool + base16 * 16, pin[b], bytes, cudaMemcpyHostToDevice, up);
          if (ce == cudaSuccess) ce = cudaEventRecord(ev[b], up);
          ++chunkNo;
        }
        t0 = t1;
      }
    }
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(up);
    cleanup();
    if (ce != cudaSuccess) return fail(BF_CUDA_ERROR_RUNTIME, std::string("value pool upload: ") + cudaGetErrorString(ce));
  }
  if (!D.ownedValues.empty()) { // the copies behind merged leaves (bf_plan.cpp step 0) have been packed: drop them
    for (DirOp &d : D.ops)
      for (const std::vector<double> &v : D.ownedValues)
        if (d.val == v.data()) { d.val = nullptr; break; }
    D.ownedValues.clear();
    D.ownedValues.shrink_to_fit();
  }
  // (a value-less plan keeps the zeros of the memset above: it is ordered on P->stream, like the synthetic fill that
  // follows.  A second cudaMemset here ran on the LEGACY stream, which a non-blocking stream does not wait for, and
  // could zero tiles the fill had already written.)
  if (D.shardWorld > 1) { // block-row shard: which ranks read which item's rows
    std::vector<uint32_t> peers;
    peers.reserve(nI);
    for (const Stage &S : D.stages) peers.insert(peers.end(), S.itemPeers.begin(), S.itemPeers.end());
    BF_CUDA_TRY(cudaMalloc(&dd.itemPeers, std::max<size_t>(nI, 1) * sizeof(uint32_t)));
    if (nI) BF_CUDA_TRY(cudaMemcpy(dd.itemPeers, peers.data(), nI * sizeof(uint32_t), cudaMemcpyHostToDevice));
  }
  // fused single-launch tables (1-RHS apply): uploaded once, host copies dropped
  FusedTables &F = D.fused;
  if (F.enabled) {
    auto up = [&](auto **dst, const auto &vec) -> cudaError_t {
      using V = typename std::remove_reference<decltype(vec)>::type::value_type;
      cudaError_t e = cudaMalloc(dst, std::max<size_t>(vec.size(), 1) * sizeof(V));
      if (e == cudaSuccess && !vec.empty()) e = cudaMemcpy(*dst, vec.data(), vec.size() * sizeof(V), cudaMemcpyHostToDevice);
      return e;
    };
    BF_CUDA_TRY(up(&dd.fItems, F.items));
    BF_CUDA_TRY(up(&dd.fItemDeps, F.itemDeps));
    BF_CUDA_TRY(up(&dd.fTiles, F.tiles));
    BF_CUDA_TRY(up(&dd.fTileDeps, F.tileDeps));
    BF_CUDA_TRY(up(&dd.fqItemBegin, F.qItemBegin));
    BF_CUDA_TRY(up(&dd.fqTileBegin, F.qTileBegin));
    dd.numCounters = F.numCounters;
    BF_CUDA_TRY(cudaMalloc(&dd.counters, std::max<size_t>(2 * (size_t)F.numCounters, 1) * sizeof(uint32_t)));
    BF_CUDA_TRY(cudaMemsetAsync(dd.counters, 0, std::max<size_t>(2 * (size_t)F.numCounters, 1) * sizeof(uint32_t), P->stream));
    dd.fusedWsRows = F.wsRows;
    dd.tableBytes += (int64_t)(F.items.size() * 24 + F.tiles.size() * 16 + 2 * (Q + 1) * 4 + 2 * (size_t)F.numCounters * 4);
    dd.fusedReady = true;
    dd.fNumItems = F.items.size(); dd.fNumTiles = F.tiles.size();
    F.items = {}; F.itemDeps = {}; F.tiles = {}; F.tileDeps = {}; // device-resident now
  }
  // The tables went up with cudaMemcpy from pageable memory: it returns once the data is STAGED, the DMA itself runs
  // on the legacy stream, which the plan's non-blocking stream (and the caller's) does not wait for.  Land everything
  // before the first apply can be enqueued.
  BF_CUDA_TRY(cudaStreamSynchronize(cudaStreamLegacy));
  BF_CUDA_TRY(cudaStreamSynchronize(P->stream));
  dd.built = true;
  return BF_CUDA_OK;
}
} // namespace bfc

extern "C" {

int32_t bfCudaAbiVersion(void) { return BF_CUDA_ABI_VERSION; }
const char *bfCudaLastErrorMessage(void) { return g_lastError.c_str(); }
const char *bfCudaStatusString(BfCudaStatus s) {
  switch (s) {
  case BF_CUDA_OK: return "ok";
  case BF_CUDA_ERROR_INVALID_ARGUMENT: return "invalid argument";
  case BF_CUDA_ERROR_NOT_IMPLEMENTED: return "not implemented";
  case BF_CUDA_ERROR_OUT_OF_MEMORY: return "out of memory";
  case BF_CUDA_ERROR_RUNTIME: return "CUDA runtime error";
  case BF_CUDA_ERROR_NO_DEVICE: return "no usable CUDA device (there is no CPU fallback)";
  case BF_CUDA_ERROR_IO: return "file error";
  case BF_CUDA_ERROR_INCOMPATIBLE_SHAPES: return "incompatible shapes";
  }
  return "unknown status";
}

BfCudaStatus bfCudaPlanDestroy(BfCudaPlan *P) {
  if (!P) return BF_CUDA_OK;
  DeviceGuard g(P->device);
  if (P->stream) cudaStreamSynchronize(P->stream);
  for (auto &d : P->dir) d.release();
  cudaFree(P->scratch); cudaFree(P->ioX); cudaFree(P->ioY);
  for (int i = 0; i < 2; ++i) {
    if (P->stage[i]) cudaFreeHost(P->stage[i]);
    if (P->stageDone[i]) cudaEventDestroy(P->stageDone[i]);
  }
  cudaFree(P->inPerm); cudaFree(P->outPerm); cudaFree(P->invInPerm); cudaFree(P->invOutPerm);
  for (cudaEvent_t e : P->events) cudaEventDestroy(e);
  if (P->stream) cudaStreamDestroy(P->stream);
  delete P;
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaPlanCreate(const BfCudaPlanDesc *desc, BfCudaPlan **out) {
  if (!desc || !out) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  *out = nullptr;
  if (desc->abiVersion != BF_CUDA_ABI_VERSION) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "ABI version mismatch");
  if (desc->dtype != BF_CUDA_REAL_F64 && desc->dtype != BF_CUDA_COMPLEX_F64)
    return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "unknown dtype");
  int nDev = 0;
  if (cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0) {
    cudaGetLastError();
    return fail(BF_CUDA_ERROR_NO_DEVICE, "no CUDA device visible; this library has no CPU fallback");
  }
  int dev = desc->device;
  if (dev < 0) BF_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= nDev) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "device ordinal out of range");
  DeviceGuard g(dev);
  if (!g.ok) return fail(BF_CUDA_ERROR_RUNTIME, "cudaSetDevice failed");
  cudaDeviceProp prop;
  BF_CUDA_TRY(cudaGetDeviceProperties(&prop, dev));
  // sm_100a cubins run on compute capability 10.x only (an "a" target has no forward compatibility: sm_120 would fail at launch)
  if (prop.major != 10) return fail(BF_CUDA_ERROR_NO_DEVICE, "device is not sm_100-class; kernels are built for sm_100a only");
  BF_CUDA_TRY(configureKernels());

  std::unique_ptr<BfCudaPlan> P(new BfCudaPlan());
  P->device = dev;
  P->numSMs = prop.multiProcessorCount;
  P->cooperativeOk = prop.cooperativeLaunch != 0;
  if (const char *e = std::getenv("BF_CUDA_FUSED")) P->fusedMode = std::atoi(e) != 0 ? 1 : 0;
  P->dtype = desc->dtype;
  P->numRows = desc->numRows;
  P->numCols = desc->numCols;
  P->numOps = desc->numOps;
  BF_CUDA_TRY(cudaStreamCreateWithFlags(&P->stream, cudaStreamNonBlocking));
  const int numQueues = P->numSMs * kWarpsPerCta;
  std::vector<int> want{BF_CUDA_OP_N};
  if (desc->buildTranspose & 1) want.push_back(BF_CUDA_OP_T);
  if ((desc->buildTranspose & 2) && desc->dtype == BF_CUDA_COMPLEX_F64) want.push_back(BF_CUDA_OP_H);
  for (int op : want) {
    std::string err;
    BfCudaStatus st = compileDirection(*desc, op, numQueues, P->dir[op].host, err);
    if (st == BF_CUDA_OK) st = uploadDirection(P.get(), P->dir[op]);
    else g_lastError = err;
    if (st != BF_CUDA_OK) { bfCudaPlanDestroy(P.release()); return st; }
    // host value pointers are only valid during this call
    for (DirOp &d : P->dir[op].host.ops) d.val = nullptr;
  }
  *out = P.release();
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaPlanGetInfo(const BfCudaPlan *P, BfCudaPlanInfo *info) {
  if (!P || !info) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  std::memset(info, 0, sizeof *info);
  const Direction &N = P->dir[0].host;
  info->dtype = P->dtype; info->device = P->device;
  info->numRows = P->numRows; info->numCols = P->numCols; info->numOps = P->numOps;
  info->numEntries = N.entries;
  info->numStagesN = (int64_t)N.stages.size();
  const DevDirection &T = P->dir[1].built ? P->dir[1] : P->dir[2];
  info->numStagesT = T.built ? (int64_t)T.host.stages.size() : 0;
  info->numItemsN = N.numItems; info->numTilesN = N.numTiles;
  for (const auto &d : P->dir) if (d.built) { info->poolBytes += (int64_t)d.host.pool16 * 16; info->tableBytes += d.tableBytes; }
  info->workspaceRowsN = N.wsRows;
  info->workspaceRowsT = T.built ? T.host.wsRows : 0;
  info->maxTileK = N.maxTileK; info->maxItemRows = N.maxItemRows;
  return BF_CUDA_OK;
}

static BfCudaStatus uploadPerm(const int64_t *perm, int64_t n, int32_t **dFwd, int32_t **dInv) {
  cudaFree(*dFwd); cudaFree(*dInv);
  *dFwd = *dInv = nullptr;
  if (!perm) return BF_CUDA_OK;
  std::vector<int32_t> f((size_t)n), inv((size_t)n, -1);
  for (int64_t i = 0; i < n; ++i) {
    if (perm[i] < 0 || perm[i] >= n || inv[(size_t)perm[i]] >= 0) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "permutation is not a bijection");
    f[(size_t)i] = (int32_t)perm[i];
    inv[(size_t)perm[i]] = (int32_t)i;
  }
  BF_CUDA_TRY(cudaMalloc(dFwd, std::max<size_t>((size_t)n, 1) * 4));
  BF_CUDA_TRY(cudaMalloc(dInv, std::max<size_t>((size_t)n, 1) * 4));
  BF_CUDA_TRY(cudaMemcpy(*dFwd, f.data(), (size_t)n * 4, cudaMemcpyHostToDevice));
  BF_CUDA_TRY(cudaMemcpy(*dInv, inv.data(), (size_t)n * 4, cudaMemcpyHostToDevice));
  BF_CUDA_TRY(cudaStreamSynchronize(cudaStreamLegacy)); // pageable copies: landed before an apply on another stream reads them
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaPlanSetPerms(BfCudaPlan *P, const int64_t *inPerm, const int64_t *outPerm) {
  if (!P) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null plan");
  DeviceGuard g(P->device);
  BF_CUDA_TRY(cudaStreamSynchronize(P->stream));
  BfCudaStatus st = uploadPerm(inPerm, P->numCols, &P->inPerm, &P->invInPerm);
  if (st != BF_CUDA_OK) return st;
  return uploadPerm(outPerm, P->numRows, &P->outPerm, &P->invOutPerm);
}

BfCudaStatus bfCudaPlanSetProfiling(BfCudaPlan *P, int32_t enable) {
  if (!P) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null plan");
  P->profiling = enable != 0;
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaPlanSetFused(BfCudaPlan *P, int32_t enable) {
  if (!P) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null plan");
  P->fusedMode = enable < 0 ? 2 : (enable != 0 ? 1 : 0);
  return BF_CUDA_OK;
}
int32_t bfCudaPlanGetFused(const BfCudaPlan *P, int32_t op) {
  if (!P || op < 0 || op > 2 || !P->dir[op].built || !P->dir[op].fusedReady || !P->cooperativeOk || P->profiling) return 0;
  return P->fusedMode == 1 || (P->fusedMode == 2 && (uint64_t)P->dir[op].host.pool16 * 16ull >= kFusedMinPoolBytes);
}

int64_t bfCudaPlanLastLaunchCount(const BfCudaPlan *P) { return P ? P->lastLaunches : 0; }
void *bfCudaPlanGetStream(const BfCudaPlan *P) { return P ? (void *)P->stream : nullptr; }

BfCudaStatus bfCudaPlanGetStageTimes(BfCudaPlan *P, float *ms, int64_t capacity, int64_t *numStages) {
  if (!P || !numStages) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  DeviceGuard g(P->device);
  *numStages = P->lastStages;
  if (!ms) return BF_CUDA_OK;
  if ((int64_t)P->events.size() < P->lastStages + 1) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "no profiled apply recorded");
  BF_CUDA_TRY(cudaEventSynchronize(P->events[(size_t)P->lastStages]));
  for (int64_t s = 0; s < P->lastStages && s < capacity; ++s)
    BF_CUDA_TRY(cudaEventElapsedTime(&ms[s], P->events[(size_t)s], P->events[(size_t)s + 1]));
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaPlanGetStageEntries(const BfCudaPlan *P, int32_t op, int64_t *entries, int64_t capacity,
                                       int64_t *numStages) {
  if (!P || !numStages || op < 0 || op > 2 || !P->dir[op].built) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "direction not built");
  const Direction &D = P->dir[op].host;
  *numStages = (int64_t)D.stages.size();
  for (int64_t s = 0; entries && s < *numStages && s < capacity; ++s) entries[s] = D.stages[(size_t)s].entries;
  return BF_CUDA_OK;
}

static BfCudaStatus ensureBytes(unsigned char **p, size_t *have, size_t want) {
  if (*have >= want) return BF_CUDA_OK;
  cudaFree(*p);
  *p = nullptr; *have = 0;
  BF_CUDA_TRY(cudaMalloc(p, want));
  *have = want;
  return BF_CUDA_OK;
}

// Stages [sBegin, sEnd) of one apply; sEnd < 0 = all.  staged = a stage range was requested: the
// RHS block must then fit one workspace chunk so that the workspace persists between calls.
static BfCudaStatus applyImpl(BfCudaPlan *P, int32_t op, const void *d_x, int64_t ldx, int64_t nrhs, void *d_y,
                              int64_t ldy, void *streamOpaque, int64_t sBegin, int64_t sEnd, bool staged) {
  if (!P) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null plan");
  if (op < 0 || op > 2 || !P->dir[op].built) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "direction not built (set buildTranspose)");
  if (nrhs < 0 || (nrhs > 0 && (ldx < nrhs || ldy < nrhs))) return fail(BF_CUDA_ERROR_INCOMPATIBLE_SHAPES, "ldx / ldy smaller than nrhs");
  if (nrhs == 0) { P->lastLaunches = 0; return BF_CUDA_OK; }
  if (!d_x || !d_y) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null operand");
  if (reinterpret_cast<uintptr_t>(d_y) & (P->dtype == BF_CUDA_COMPLEX_F64 ? 15 : 7))
    return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "d_y must be aligned to its element size");
  if (reinterpret_cast<uintptr_t>(d_x) & 7) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "d_x must be 8-byte aligned");
  DeviceGuard g(P->device);
  DevDirection &dd = P->dir[op];
  const Direction &D = dd.host;
  cudaStream_t stream = streamOpaque ? (cudaStream_t)streamOpaque : P->stream;
  const bool cplx = P->dtype == BF_CUDA_COMPLEX_F64;
  const size_t es = cplx ? 16 : 8;
  const int ed = cplx ? 2 : 1;
  const int32_t *scatterIn = op == BF_CUDA_OP_N ? P->inPerm : P->invOutPerm;
  const int32_t *scatterOut = op == BF_CUDA_OP_N ? P->outPerm : P->invInPerm;

  // 1 RHS, whole apply: ONE cooperative launch over the fused tables (flat workspace, readiness counters)
  // Measured A/B (tools/fused_ab.py, profiles/r2c_fused_ab.jsonl): 10.87 -> 10.66 ms at N = 524 288 (0.942 -> 0.961
  // of the HBM peak), but 0.591 -> 0.613 ms at 65 536 and 0.152 -> 0.189 ms at 16 384, where a stage is a handful of
  // tiles per warp and the hand-over through counters costs more than the PDL-chained launches; hence the size
  // switch (fusedMode: 0 never, 1 always, 2 = operators whose pool is at least kFusedMinPoolBytes).
  const bool fusedWanted = P->fusedMode == 1 || (P->fusedMode == 2 && (uint64_t)D.pool16 * 16ull >= kFusedMinPoolBytes);
  const bool fused = nrhs == 1 && dd.fusedReady && P->cooperativeOk && fusedWanted && !P->profiling && !staged && D.shardWorld <= 1;
  // column chunking keeps the workspace within budget
  const size_t rowsPad = (size_t)(fused ? std::max(D.wsRows, dd.fusedWsRows) : D.wsRows) + 2;
  const size_t xRowsPad = (size_t)D.rowsIn + 2;
  int64_t chunk = nrhs;
  // the budget: what the device has free (plus what the plan already holds for this purpose) minus a margin,
  // so a 256-column block runs as ONE chunk whenever the memory is there — every extra chunk re-reads the pool
  size_t budget = P->maxScratchBytes;
  // (only when a block of columns does not fit what the plan already holds: cudaMemGetInfo stalls while the GPU is
  // busy, it must not sit on the path of a 1-column apply)
  if (nrhs > 1 && (rowsPad + xRowsPad) * (size_t)nrhs * es > P->scratchBytes) {
    size_t freeB = 0, totalB = 0;
    if (cudaMemGetInfo(&freeB, &totalB) == cudaSuccess) {
      const size_t margin = size_t(2) << 30;
      const size_t avail = freeB + P->scratchBytes > margin ? freeB + P->scratchBytes - margin : 0;
      budget = std::min(budget, avail);
    }
  }
  while (chunk > 1 && (rowsPad + xRowsPad) * (size_t)chunk * es + 512 > budget) chunk = (chunk + 1) / 2;
  const bool sharded = D.shardWorld > 1;
  if (sharded) {
    // one rank of a block-row sharded operator: peers store into the workspace, so it is the externally owned,
    // peer-mapped region, and the hand-over protocol (bf_kernels.cu) is written for one column per apply
    if (!P->shardDev || !P->extWs) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "a block-row shard is applied through bfCudaShardRankApply / bfCudaShardedApply* (its peers must be connected)");
    if (nrhs != 1) return fail(BF_CUDA_ERROR_NOT_IMPLEMENTED, "a block-row shard applies one column per call");
    if (rowsPad * es > P->extWsBytes) return fail(BF_CUDA_ERROR_RUNTIME, "exchange workspace smaller than the plan's workspace");
    chunk = 1;
  }
  const size_t wsBytes = sharded ? 0 : ((rowsPad * (size_t)chunk * es + 255) / 256) * 256;
  const size_t need = wsBytes + xRowsPad * (size_t)chunk * es + 256;
  if (P->scratchBytes < need) {
    BF_CUDA_TRY(cudaStreamSynchronize(stream));
    BfCudaStatus st = ensureBytes(&P->scratch, &P->scratchBytes, need);
    if (st != BF_CUDA_OK) return st;
  }
  const size_t nS = D.stages.size();
  const size_t Q = (size_t)D.numQueues;
  if (sEnd < 0) sEnd = (int64_t)nS;
  if (sBegin < 0 || sEnd > (int64_t)nS || sBegin > sEnd) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "stage range");
  if (staged && chunk != nrhs) return fail(BF_CUDA_ERROR_NOT_IMPLEMENTED, "stage-by-stage apply needs the RHS block to fit one workspace chunk");
  if (P->profiling) {
    while (P->events.size() < nS + 1) { cudaEvent_t e; BF_CUDA_TRY(cudaEventCreate(&e)); P->events.push_back(e); }
  }
  int64_t launches = 0;
  for (int64_t c0 = 0; c0 < nrhs; c0 += chunk) {
    const int64_t nc = std::min(chunk, nrhs - c0);
    const unsigned char *xp = static_cast<const unsigned char *>(d_x) + (size_t)c0 * es;
    unsigned char *yp = static_cast<unsigned char *>(d_y) + (size_t)c0 * es;
    StageLaunch a{};
    a.pool = dd.pool;
    a.base[kSpaceY] = yp; a.ld[kSpaceY] = ldy;
    a.base[kSpaceW] = sharded ? P->extWs : P->scratch; a.ld[kSpaceW] = nc;
    a.outPerm = scatterOut;
    a.numQueues = D.numQueues; a.nrhs = (int)nc; a.colBegin = 0;
    // overlap a stage's prologue with the previous stage's tail where stages are short (bf_kernels.cu)
    a.pdl = (uint64_t)D.pool16 * 16ull < (4ull << 30) ? 1 : 0;
    // The 1-RHS kernel bulk-copies operand segments: they must be dense and 16-byte aligned (real
    // segments may start 8 bytes into a 16-byte window, so they always go through the padded copy).
    const bool needStage = scatterIn != nullptr || (cplx && (reinterpret_cast<uintptr_t>(xp) & 15)) || (nc == 1 && (!cplx || ldx != 1));
    if (needStage && sBegin > 0) {
      a.base[kSpaceX] = P->scratch + wsBytes; a.ld[kSpaceX] = nc; // staged by the call that ran stage 0
    } else if (needStage) {
      unsigned char *xs = P->scratch + wsBytes;
      cudaError_t e = launchPermuteRows(reinterpret_cast<const double *>(xp), ldx, reinterpret_cast<double *>(xs), nc,
                                        scatterIn, D.rowsIn, (int)nc, ed, stream);
      if (e != cudaSuccess) return fail(BF_CUDA_ERROR_RUNTIME, std::string("permute launch: ") + cudaGetErrorString(e));
      ++launches;
      a.base[kSpaceX] = xs; a.ld[kSpaceX] = nc;
    } else {
      a.base[kSpaceX] = const_cast<unsigned char *>(xp); a.ld[kSpaceX] = ldx;
    }
    if (fused) {
      a.pdl = 0;
      a.items = dd.fItems; a.tiles = dd.fTiles;
      a.itemDeps = dd.fItemDeps; a.tileDeps = dd.fTileDeps;
      a.qItemBegin = dd.fqItemBegin; a.qTileBegin = dd.fqTileBegin; a.qPool16 = nullptr;
      a.numCounters = dd.numCounters;
      a.cnt = dd.counters + (size_t)(dd.applyParity & 1u) * dd.numCounters;
      a.cntNext = dd.counters + (size_t)((dd.applyParity & 1u) ^ 1u) * dd.numCounters;
      cudaError_t e = launchGemvFused(P->dtype, a, stream);
      if (e == cudaErrorCooperativeLaunchTooLarge) { // the CTAs cannot all be resident: per-stage launches from now on
        cudaGetLastError();
        P->cooperativeOk = false;
        return applyImpl(P, op, d_x, ldx, nrhs, d_y, ldy, streamOpaque, sBegin, sEnd, staged);
      }
      if (e != cudaSuccess) return fail(BF_CUDA_ERROR_RUNTIME, std::string("fused apply launch: ") + cudaGetErrorString(e));
      ++dd.applyParity;
      ++launches;
      continue;
    }
    for (size_t s = (size_t)sBegin; s < (size_t)sEnd; ++s) {
      if (P->profiling && c0 == 0) BF_CUDA_TRY(cudaEventRecord(P->events[s], stream));
      a.items = dd.items + dd.itemOff[s];
      a.tiles = dd.tiles + dd.tileOff[s];
      a.qItemBegin = dd.qItemBegin + s * (Q + 1);
      a.qTileBegin = dd.qTileBegin + s * (Q + 1);
      a.qPool16 = dd.qPool16 + s * (Q + 1);
      if (sharded) {
        a.shard = P->shardDev;
        a.itemPeers = dd.itemPeers + dd.itemOff[s];
        a.stage = (int)s;
        a.epoch = P->shardEpoch;
      }
      cudaError_t e = sharded ? launchGemvShardStage(P->dtype, a, stream)
                              : (nc == 1 ? launchGemvStage(P->dtype, a, P->numSMs, stream) : launchWideStage(P->dtype, a, P->numSMs, stream));
      if (e != cudaSuccess) return fail(BF_CUDA_ERROR_RUNTIME, std::string("stage launch: ") + cudaGetErrorString(e));
      ++launches;
    }
    if (P->profiling && c0 == 0 && sEnd == (int64_t)nS) BF_CUDA_TRY(cudaEventRecord(P->events[nS], stream));
  }
  P->lastStages = (int64_t)nS;
  P->lastLaunches = launches;
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaApplyDevice(BfCudaPlan *P, int32_t op, const void *d_x, int64_t ldx, int64_t nrhs,
                               void *d_y, int64_t ldy, void *streamOpaque) {
  return applyImpl(P, op, d_x, ldx, nrhs, d_y, ldy, streamOpaque, 0, -1, false);
}

BfCudaStatus bfCudaApplyStages(BfCudaPlan *P, int32_t op, const void *d_x, int64_t ldx, int64_t nrhs, void *d_y,
                               int64_t ldy, int64_t stageBegin, int64_t stageEnd, void *streamOpaque) {
  return applyImpl(P, op, d_x, ldx, nrhs, d_y, ldy, streamOpaque, stageBegin, stageEnd, true);
}

BfCudaStatus bfCudaPlanGetWorkspaceLayout(const BfCudaPlan *P, int32_t op, int64_t *a0, int64_t *a1, int64_t *ded) {
  if (!P || op < 0 || op > 2 || !P->dir[op].built) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "direction not built");
  const Direction &D = P->dir[op].host;
  if (a0) *a0 = D.wsArena[0];
  if (a1) *a1 = D.wsArena[1];
  if (ded) *ded = D.wsRows - D.wsArena[0] - D.wsArena[1];
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaPlanGetWorkspace(BfCudaPlan *P, int32_t op, int64_t nrhs, void **d_ws, int64_t *ld) {
  if (!P || op < 0 || op > 2 || !P->dir[op].built || nrhs <= 0 || !d_ws) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "bad argument");
  DeviceGuard g(P->device);
  const Direction &D = P->dir[op].host;
  const size_t es = P->dtype == BF_CUDA_COMPLEX_F64 ? 16 : 8;
  const size_t rowsPad = (size_t)D.wsRows + 2, xRowsPad = (size_t)D.rowsIn + 2;
  if ((rowsPad + xRowsPad) * (size_t)nrhs * es > P->maxScratchBytes) return fail(BF_CUDA_ERROR_NOT_IMPLEMENTED, "RHS block does not fit one workspace chunk");
  const size_t wsBytes = ((rowsPad * (size_t)nrhs * es + 255) / 256) * 256;
  const size_t need = wsBytes + xRowsPad * (size_t)nrhs * es + 256;
  if (P->scratchBytes < need) {
    BF_CUDA_TRY(cudaStreamSynchronize(P->stream));
    BfCudaStatus st = ensureBytes(&P->scratch, &P->scratchBytes, need);
    if (st != BF_CUDA_OK) return st;
  }
  *d_ws = P->scratch;
  if (ld) *ld = nrhs;
  return BF_CUDA_OK;
}

// ---- host-buffer apply ----------------------------------------------------------------------------
// The reference hands us malloc'ed (pageable) BfMatDenseComplex::data / BfVecReal::data.  A
// cudaMemcpyAsync from pageable memory is staged by the driver in small synchronous pieces; we stage it
// ourselves instead: the operand is gathered (any ld) into one of two pinned buffers by the host while
// the DMA engine moves the other one, and the result comes back the same way.  Buffers that are already
// pinned (bench.py's e2e leg, a caller that registered its arrays) are copied directly.
static bool hostPointerIsPinned(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

static BfCudaStatus ensureStage(BfCudaPlan *P) {
  if (P->stage[0]) return BF_CUDA_OK;
  for (int i = 0; i < 2; ++i) {
    BF_CUDA_TRY(cudaHostAlloc(&P->stage[i], kStageBytes, cudaHostAllocDefault));
    BF_CUDA_TRY(cudaEventCreateWithFlags(&P->stageDone[i], cudaEventDisableTiming));
  }
  return BF_CUDA_OK;
}

// rows x rowBytes block, host row stride hostPitch  <->  dense device block
static BfCudaStatus stagedCopy(BfCudaPlan *P, unsigned char *dev, unsigned char *host, size_t hostPitch, size_t rowBytes,
                               size_t rows, bool toDevice) {
  if (rows == 0 || rowBytes == 0) return BF_CUDA_OK;
  const bool dense = hostPitch == rowBytes;
  if (hostPointerIsPinned(host)) {
    if (dense) BF_CUDA_TRY(cudaMemcpyAsync(toDevice ? (void *)dev : (void *)host, toDevice ? (void *)host : (void *)dev, rows * rowBytes,
                                           toDevice ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, P->stream));
    else if (toDevice) BF_CUDA_TRY(cudaMemcpy2DAsync(dev, rowBytes, host, hostPitch, rowBytes, rows, cudaMemcpyHostToDevice, P->stream));
    else BF_CUDA_TRY(cudaMemcpy2DAsync(host, hostPitch, dev, rowBytes, rowBytes, rows, cudaMemcpyDeviceToHost, P->stream));
    return BF_CUDA_OK;
  }
  BfCudaStatus st = ensureStage(P);
  if (st != BF_CUDA_OK) return st;
  const size_t rowsPerChunk = std::max<size_t>(1, kStageBytes / rowBytes);
  if (rowBytes > kStageBytes) { // absurdly wide rows: let the driver stage them
    if (toDevice) BF_CUDA_TRY(cudaMemcpy2DAsync(dev, rowBytes, host, hostPitch, rowBytes, rows, cudaMemcpyHostToDevice, P->stream));
    else BF_CUDA_TRY(cudaMemcpy2DAsync(host, hostPitch, dev, rowBytes, rowBytes, rows, cudaMemcpyDeviceToHost, P->stream));
    return BF_CUDA_OK;
  }
  const size_t nChunks = (rows + rowsPerChunk - 1) / rowsPerChunk;
  auto gather = [&](unsigned char *stage, size_t r0, size_t nr, bool in) {
    if (dense) {
      if (in) std::memcpy(stage, host + r0 * rowBytes, nr * rowBytes);
      else std::memcpy(host + r0 * rowBytes, stage, nr * rowBytes);
      return;
    }
    for (size_t r = 0; r < nr; ++r) {
      if (in) std::memcpy(stage + r * rowBytes, host + (r0 + r) * hostPitch, rowBytes);
      else std::memcpy(host + (r0 + r) * hostPitch, stage + r * rowBytes, rowBytes);
    }
  };
  if (toDevice) {
    for (size_t c = 0; c < nChunks; ++c) {
      const int b = (int)(c & 1);
      const size_t r0 = c * rowsPerChunk, nr = std::min(rowsPerChunk, rows - r0);
      if (c >= 2) BF_CUDA_TRY(cudaEventSynchronize(P->stageDone[b])); // the DMA out of this buffer has finished
      gather(P->stage[b], r0, nr, true);
      BF_CUDA_TRY(cudaMemcpyAsync(dev + r0 * rowBytes, P->stage[b], nr * rowBytes, cudaMemcpyHostToDevice, P->stream));
      BF_CUDA_TRY(cudaEventRecord(P->stageDone[b], P->stream));
    }
  } else {
    // chunk c + 1 is in flight while the host scatters chunk c
    auto issue = [&](size_t c) -> cudaError_t {
      const int b = (int)(c & 1);
      const size_t r0 = c * rowsPerChunk, nr = std::min(rowsPerChunk, rows - r0);
      cudaError_t e = cudaMemcpyAsync(P->stage[b], dev + r0 * rowBytes, nr * rowBytes, cudaMemcpyDeviceToHost, P->stream);
      return e == cudaSuccess ? cudaEventRecord(P->stageDone[b], P->stream) : e;
    };
    BF_CUDA_TRY(issue(0));
    for (size_t c = 0; c < nChunks; ++c) {
      const int b = (int)(c & 1);
      const size_t r0 = c * rowsPerChunk, nr = std::min(rowsPerChunk, rows - r0);
      if (c + 1 < nChunks) BF_CUDA_TRY(issue(c + 1));
      BF_CUDA_TRY(cudaEventSynchronize(P->stageDone[b]));
      gather(P->stage[b], r0, nr, false);
    }
  }
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaApplyHost(BfCudaPlan *P, int32_t op, const void *x, int64_t ldx, int64_t nrhs, void *y,
                             int64_t ldy) {
  if (!P) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null plan");
  if (op < 0 || op > 2 || !P->dir[op].built) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "direction not built (set buildTranspose)");
  if (nrhs < 0 || (nrhs > 0 && (ldx < nrhs || ldy < nrhs))) return fail(BF_CUDA_ERROR_INCOMPATIBLE_SHAPES, "ldx / ldy smaller than nrhs");
  if (nrhs == 0) return BF_CUDA_OK;
  if (!x || !y) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null operand");
  DeviceGuard g(P->device);
  const Direction &D = P->dir[op].host;
  const size_t es = P->dtype == BF_CUDA_COMPLEX_F64 ? 16 : 8;
  const size_t rowBytes = (size_t)nrhs * es;
  BfCudaStatus st = ensureBytes(&P->ioX, &P->ioXBytes, std::max<size_t>((size_t)D.rowsIn * rowBytes, 16));
  if (st != BF_CUDA_OK) return st;
  st = ensureBytes(&P->ioY, &P->ioYBytes, std::max<size_t>((size_t)D.rowsOut * rowBytes, 16));
  if (st != BF_CUDA_OK) return st;
  // the device copies are always dense (ld == nrhs): strided operands (BfVec views, padded row strides)
  // are gathered on the host side of the copy
  st = stagedCopy(P, P->ioX, const_cast<unsigned char *>(static_cast<const unsigned char *>(x)), (size_t)ldx * es, rowBytes,
                  (size_t)D.rowsIn, true);
  if (st != BF_CUDA_OK) return st;
  st = bfCudaApplyDevice(P, op, P->ioX, nrhs, nrhs, P->ioY, nrhs, P->stream);
  if (st != BF_CUDA_OK) return st;
  st = stagedCopy(P, P->ioY, static_cast<unsigned char *>(y), (size_t)ldy * es, rowBytes, (size_t)D.rowsOut, false);
  if (st != BF_CUDA_OK) return st;
  BF_CUDA_TRY(cudaStreamSynchronize(P->stream));
  return BF_CUDA_OK;
}

BfCudaStatus bfCudaPlanFillSynthetic(BfCudaPlan *P, uint64_t seed) {
  if (!P) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null plan");
  if (P->loadedFromFile) return fail(BF_CUDA_ERROR_NOT_IMPLEMENTED, "a plan loaded from a file holds no per-leaf data: its values cannot be re-filled");
  DeviceGuard g(P->device);
  for (auto &dd : P->dir) {
    if (!dd.built) continue;
    const Direction &D = dd.host;
    for (const Stage &S : D.stages) {
      const size_t nT = S.tiles.size();
      if (!nT) continue;
      std::vector<FillPiece> fp(nT);
      for (size_t t = 0; t < nT; ++t) {
        const DirOp &d = D.ops[(size_t)S.pieces[t].op];
        FillPiece &f = fp[t];
        f.pool16 = S.tilePool16[t];
        f.rawIndex = d.rawIndex;
        f.row0 = S.pieces[t].row0; f.col0 = S.pieces[t].col0;
        f.storedCols = (int32_t)d.storedCols;
        f.rows = S.tiles[t].rows; f.kc = S.tiles[t].kc;
        f.swapRC = d.swapRC; f.conj = d.conj; f.kind = (uint8_t)d.kind; f.pad = 0;
        // scale by the contraction length of the leaf as it acts in the forward direction
        f.scale = syntheticScale(D.op == BF_CUDA_OP_N ? d.k : d.m);
      }
      FillPiece *dp = nullptr;
      BF_CUDA_TRY(cudaMalloc(&dp, nT * sizeof(FillPiece)));
      // (stream-ordered with the fill kernel; a plain cudaMemcpy from pageable memory returns before its DMA has
      // landed and the plan's stream is non-blocking)
      cudaError_t e = cudaMemcpyAsync(dp, fp.data(), nT * sizeof(FillPiece), cudaMemcpyHostToDevice, P->stream);
      if (e == cudaSuccess) e = launchFillSynthetic(P->dtype, dd.pool, dp, (long long)nT, seed, P->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(P->stream);
      cudaFree(dp);
      if (e != cudaSuccess) return fail(BF_CUDA_ERROR_RUNTIME, std::string("synthetic fill: ") + cudaGetErrorString(e));
    }
  }
  return BF_CUDA_OK;
}

double bfCudaSyntheticValue(uint64_t seed, int64_t opIndex, int64_t storedLinear, int32_t part, int64_t k) {
  return syntheticValue(seed, opIndex, storedLinear, part, syntheticScale(k));
}
double bfCudaSyntheticDiagValue(uint64_t seed, int64_t opIndex, int64_t i) {
  return 1.0 + 0.5 * syntheticValue(seed, opIndex, i, 0, 1.0);
}

} // extern "C"
