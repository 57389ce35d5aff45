// bf_internal.hpp — definitions shared by the translation units that implement the C ABI
// (bf_capi.cu: plan life cycle and apply; bf_helm2.cu: on-device construction of factor values).
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "bf_cuda.h"
#include "bf_kernels.cuh"
#include "bf_plan.hpp"

namespace bfc {
// Records the thread-local message behind bfCudaLastErrorMessage and returns `st`.
BfCudaStatus fail(BfCudaStatus st, const std::string &msg);
} // namespace bfc

#define BF_CUDA_TRY(expr)                                                                     \
  do {                                                                                        \
    cudaError_t e_ = (expr);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      BfCudaStatus st_ = e_ == cudaErrorMemoryAllocation ? BF_CUDA_ERROR_OUT_OF_MEMORY       \
                                                          : BF_CUDA_ERROR_RUNTIME;            \
      return bfc::fail(st_, std::string(#expr) + ": " + cudaGetErrorString(e_));             \
    }                                                                                         \
  } while (0)

namespace bfc {
struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) return;
    ok = prev == dev || cudaSetDevice(dev) == cudaSuccess;
    if (prev == dev) prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

struct DevDirection {
  bool built = false;
  Direction host;
  // all stages concatenated; per-stage element offsets below
  DevItem *items = nullptr;
  DevTile *tiles = nullptr;
  uint32_t *qItemBegin = nullptr, *qTileBegin = nullptr;
  uint64_t *qPool16 = nullptr;
  unsigned char *pool = nullptr;
  std::vector<size_t> itemOff, tileOff; // per stage
  int64_t tableBytes = 0;
  // fused single-launch tables (FusedTables) and the two alternating counter sets
  bool fusedReady = false;
  DevItem *fItems = nullptr;
  DevTile *fTiles = nullptr;
  ItemDep *fItemDeps = nullptr;
  TileDep *fTileDeps = nullptr;
  uint32_t *fqItemBegin = nullptr, *fqTileBegin = nullptr;
  size_t fNumItems = 0, fNumTiles = 0;
  uint32_t *itemPeers = nullptr; // block-row shards: per item (all stages concatenated like `items`), ranks that read its rows
  uint32_t *counters = nullptr; // [2][numCounters]
  uint32_t numCounters = 0;
  int64_t fusedWsRows = 0;
  unsigned applyParity = 0;
  void release() {
    cudaFree(items); cudaFree(tiles); cudaFree(qItemBegin); cudaFree(qTileBegin); cudaFree(qPool16); cudaFree(pool);
    cudaFree(fItems); cudaFree(fTiles); cudaFree(fItemDeps); cudaFree(fTileDeps); cudaFree(fqItemBegin); cudaFree(fqTileBegin);
    cudaFree(counters); cudaFree(itemPeers);
    itemPeers = nullptr;
    items = nullptr; tiles = nullptr; qItemBegin = qTileBegin = nullptr; qPool16 = nullptr; pool = nullptr;
    fItems = nullptr; fTiles = nullptr; fItemDeps = nullptr; fTileDeps = nullptr; fqItemBegin = fqTileBegin = nullptr;
    counters = nullptr;
    built = false; fusedReady = false;
  }
};
} // namespace bfc

constexpr size_t kStageBytes = size_t(8) << 20; // one pinned staging buffer
constexpr uint64_t kFusedMinPoolBytes = 24ull << 30; // single-launch 1-RHS apply from this pool size on (bf_capi.cu applyImpl)

namespace bfc {
// Device tables + pool of one compiled direction (bf_capi.cu); values of leaves with host pointers are packed and
// uploaded, the pool of a value-less direction is zeroed.
BfCudaStatus uploadDirection(struct ::BfCudaPlan *P, DevDirection &dd);
} // namespace bfc

struct BfCudaPlan {
  int device = 0;
  int numSMs = 148;
  int dtype = BF_CUDA_REAL_F64;
  int64_t numRows = 0, numCols = 0, numOps = 0;
  bfc::DevDirection dir[3];
  cudaStream_t stream = nullptr;
  // scratch: [workspace | staged X], (re)grown on demand
  unsigned char *scratch = nullptr;
  size_t scratchBytes = 0;
  // host-API I/O buffers
  unsigned char *ioX = nullptr, *ioY = nullptr;
  size_t ioXBytes = 0, ioYBytes = 0;
  // pinned double buffer between pageable caller memory and ioX / ioY (bfCudaApplyHost)
  unsigned char *stage[2] = {nullptr, nullptr};
  cudaEvent_t stageDone[2] = {nullptr, nullptr};
  // fused caller-side permutations (scatter form) and their inverses
  int32_t *inPerm = nullptr, *outPerm = nullptr, *invInPerm = nullptr, *invOutPerm = nullptr;
  int64_t lastLaunches = 0;
  bool profiling = false;
  std::vector<cudaEvent_t> events;
  int64_t lastStages = 0;
  size_t maxScratchBytes = size_t(96) << 30; // upper bound of the workspace; the real budget is the free memory (applyImpl)
  bool cooperativeOk = false; // device supports cooperative launches (the fused 1-RHS apply needs co-resident CTAs)
  // block-row shard (bf_comm.cu): the workspace lives in memory the peers can store into, and every stage kernel
  // gets the hand-over context; one column per apply, `shardEpoch` counts them
  unsigned char *extWs = nullptr;      // not owned
  size_t extWsBytes = 0;
  const bfc::ShardLaunch *shardDev = nullptr; // device copy, not owned
  uint32_t shardEpoch = 0;
  bool loadedFromFile = false; // bfCudaPlanLoad: tables and pool only, no per-leaf host data (no value (re)build)
  int fusedMode = 2;          // bfCudaPlanSetFused / BF_CUDA_FUSED: 0 per-stage launches, 1 single launch, 2 by size
};

