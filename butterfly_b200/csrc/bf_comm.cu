// bf_comm.cu — one operator over several GPUs behind the C ABI (SURVEY.md §8b "explicit ABI", §8e).
//
// The reference is single-process, single-thread C (src/linalg.c:155 calls bfMatMul(A, V[j]) in a loop); nothing in
// it shards an operator.  Two modes:
//   RHS columns   the plan is replicated, device g applies its slice of the columns; no data crosses GPUs.
//   block rows    ONE operator, every device stores 1 / world of the factors (bf_shard.cpp decides who applies
//                 which leaf).  The rows of an intermediate level that another device reads are stored straight
//                 into that device's workspace by the producing stage kernel (peer memory over NVLink) and
//                 handed over with one flag per (stage, peer) — bf_kernels.cu, MODE 2.  No NCCL, no collective,
//                 no host round trip between stages: all stages of all devices are enqueued at once.
// Layers, bottom up:
//   BfCudaShardDescs  the per-rank descriptors of a block-row sharded operator (host only)
//   BfCudaShardRank   one rank at run time: its plan, its peer-visible workspace and flags.  Peers are connected
//                     through BfCudaShardHandle: raw pointers when they live in the same process, CUDA IPC handles
//                     when every GPU has a process of its own (bench.py under torchrun)
//   BfCudaComm / BfCudaShardedPlan   all ranks in ONE process, driven by one host thread — what the reference's
//                     callers (bfMatMul through bfCudaMatAttachSharded, bfSolveGMRES) use
#include <cuda_runtime.h>
#include <unistd.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "bf_internal.hpp"
#include "bf_shard.hpp"

using namespace bfc;

static_assert(kMaxRanks >= kMaxShardRanks, "rank limits");

// ---- per-rank descriptors --------------------------------------------------------------------------------
struct BfCudaShardDescs {
  ShardedOps S;
  std::vector<int64_t> bufRows;
  std::vector<std::vector<BfCudaRawOp>> rankOps; // [world]
  std::vector<BfCudaPlanDesc> descs;             // [world]
};

extern "C" BfCudaStatus bfCudaShardDescsCreate(const BfCudaPlanDesc *desc, int32_t world, const double *fractions,
                                               BfCudaShardDescs **out) {
  if (!desc || !out) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  *out = nullptr;
  if (desc->abiVersion != BF_CUDA_ABI_VERSION) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "ABI version mismatch");
  std::unique_ptr<BfCudaShardDescs> sd(new BfCudaShardDescs());
  std::string err;
  BfCudaStatus st = shardOps(*desc, world, fractions, sd->S, err);
  if (st != BF_CUDA_OK) return fail(st, err);
  sd->bufRows.assign(desc->bufRows, desc->bufRows + desc->numBufs);
  sd->rankOps.resize((size_t)world);
  sd->descs.resize((size_t)world);
  for (int r = 0; r < world; ++r) {
    maskForRank(sd->S, r, sd->rankOps[(size_t)r]);
    BfCudaPlanDesc &d = sd->descs[(size_t)r];
    d = *desc;
    d.bufRows = sd->bufRows.data();
    d.numOps = (int64_t)sd->rankOps[(size_t)r].size();
    d.ops = sd->rankOps[(size_t)r].data();
    d.buildTranspose = 0;
    d.shardWorld = world;
    d.shardRank = r;
    if (world == 1) { // nothing to shard: an ordinary descriptor (owner bits cleared)
      d.shardWorld = 0;
      for (BfCudaRawOp &o : sd->rankOps[0]) o.flags &= ~(0xFFu << BF_CUDA_LEAF_OWNER_SHIFT);
    }
  }
  *out = sd.release();
  return BF_CUDA_OK;
}

extern "C" const BfCudaPlanDesc *bfCudaShardDescsGet(const BfCudaShardDescs *sd, int32_t rank) {
  return sd && rank >= 0 && rank < sd->S.world ? &sd->descs[(size_t)rank] : nullptr;
}

extern "C" BfCudaStatus bfCudaShardDescsGetLayout(const BfCudaShardDescs *sd, int64_t *numOps, int64_t *yCuts, int32_t *owner,
                                                  int64_t *originOp, int64_t *originRow) {
  if (!sd) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  const size_t n = sd->S.ops.size();
  if (numOps) *numOps = (int64_t)n;
  if (yCuts) std::copy(sd->S.yCuts.begin(), sd->S.yCuts.end(), yCuts);
  if (owner) std::copy(sd->S.owner.begin(), sd->S.owner.end(), owner);
  if (originOp) std::copy(sd->S.originOp.begin(), sd->S.originOp.end(), originOp);
  if (originRow) std::copy(sd->S.originRow.begin(), sd->S.originRow.end(), originRow);
  return BF_CUDA_OK;
}

extern "C" void bfCudaShardDescsFree(BfCudaShardDescs *sd) { delete sd; }

// ---- one rank at run time -----------------------------------------------------------------------------------
struct BfCudaShardRank {
  BfCudaPlan *plan = nullptr;
  int device = 0, world = 1, rank = 0, numStages = 0;
  unsigned char *ws = nullptr;   // peer-visible workspace (cudaMalloc: IPC-exportable)
  size_t wsBytes = 0;
  uint32_t *flags = nullptr;     // [numStages][world] written by peers
  uint32_t *done = nullptr;      // [numStages]
  ShardLaunch host{};
  ShardLaunch *dev = nullptr;
  void *opened[kMaxRanks][2] = {}; // IPC mappings to close
  bool connected = false;
  int64_t sentRows = 0;            // workspace rows this rank pushes to peers per apply
};

extern "C" BfCudaStatus bfCudaShardRankDestroy(BfCudaShardRank *R) {
  if (!R) return BF_CUDA_OK;
  DeviceGuard g(R->device);
  if (R->plan) {
    cudaStreamSynchronize(R->plan->stream);
    bfCudaPlanDestroy(R->plan);
  }
  cudaDeviceSynchronize(); // nothing of ours may still be running when the peer mappings go away
  for (auto &o : R->opened)
    for (void *p : o)
      if (p) cudaIpcCloseMemHandle(p);
  cudaFree(R->ws); cudaFree(R->flags); cudaFree(R->done); cudaFree(R->dev);
  delete R;
  return BF_CUDA_OK;
}

// Takes ownership of `plan` (destroyed on failure too).
static BfCudaStatus wrapRankPlan(BfCudaPlan *plan, BfCudaShardRank **out) {
  std::unique_ptr<BfCudaShardRank> R(new BfCudaShardRank());
  R->plan = plan;
  BfCudaPlan *P = R->plan;
  R->device = P->device;
  const Direction &D = P->dir[0].host;
  R->world = std::max(D.shardWorld, 1);
  R->rank = D.shardRank;
  R->numStages = (int)D.stages.size();
  if (R->world == 1) { *out = R.release(); return BF_CUDA_OK; } // an ordinary plan behind the same interface
  if (R->numStages > 32) { bfCudaShardRankDestroy(R.release()); return fail(BF_CUDA_ERROR_NOT_IMPLEMENTED, "more than 32 stages in a sharded plan"); }
  DeviceGuard g(R->device);
  const size_t es = P->dtype == BF_CUDA_COMPLEX_F64 ? 16 : 8;
  R->wsBytes = (((size_t)D.wsRows + 2) * es + 255) / 256 * 256;
  auto bail = [&](cudaError_t e, const char *what) {
    bfCudaShardRankDestroy(R.release());
    return fail(e == cudaErrorMemoryAllocation ? BF_CUDA_ERROR_OUT_OF_MEMORY : BF_CUDA_ERROR_RUNTIME, std::string(what) + ": " + cudaGetErrorString(e));
  };
  cudaError_t e;
  if ((e = cudaMalloc(&R->ws, R->wsBytes)) != cudaSuccess) return bail(e, "workspace");
  if ((e = cudaMemset(R->ws, 0, R->wsBytes)) != cudaSuccess) return bail(e, "workspace");
  const size_t nFlags = (size_t)R->numStages * (size_t)R->world;
  if ((e = cudaMalloc(&R->flags, nFlags * 4)) != cudaSuccess) return bail(e, "flags");
  if ((e = cudaMemset(R->flags, 0, nFlags * 4)) != cudaSuccess) return bail(e, "flags");
  if ((e = cudaMalloc(&R->done, (size_t)R->numStages * 4)) != cudaSuccess) return bail(e, "done counters");
  if ((e = cudaMemset(R->done, 0, (size_t)R->numStages * 4)) != cudaSuccess) return bail(e, "done counters");
  if ((e = cudaMalloc(&R->dev, sizeof(ShardLaunch))) != cudaSuccess) return bail(e, "launch context");
  // the memsets above run on the legacy stream; the stage kernels (and peers) use non-blocking streams
  if ((e = cudaStreamSynchronize(cudaStreamLegacy)) != cudaSuccess) return bail(e, "workspace clear");
  ShardLaunch &h = R->host;
  h.world = R->world; h.rank = R->rank; h.numStages = R->numStages;
  h.done = R->done;
  h.totalWarps = (uint32_t)((D.numQueues + kWarpsPerCta - 1) / kWarpsPerCta) * (uint32_t)kWarpsPerCta;
  for (int s = 0; s < R->numStages; ++s) {
    h.recvMask[s] = D.stages[(size_t)s].recvMask;
    h.sendMask[s] = D.stages[(size_t)s].sendMask;
    R->sentRows += D.stages[(size_t)s].sentRows;
  }
  *out = R.release();
  return BF_CUDA_OK;
}

extern "C" BfCudaStatus bfCudaShardRankCreate(const BfCudaPlanDesc *rankDesc, BfCudaShardRank **out) {
  if (!rankDesc || !out) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  *out = nullptr;
  BfCudaPlan *plan = nullptr;
  BfCudaStatus st = bfCudaPlanCreate(rankDesc, &plan);
  return st == BF_CUDA_OK ? wrapRankPlan(plan, out) : st;
}

// One rank from the file its plan was saved to (bfCudaPlanSave(bfCudaShardRankGetPlan(rank), path)): the factors
// of a sharded operator are built once per rank and re-read afterwards.
extern "C" BfCudaStatus bfCudaShardRankLoad(const char *path, int32_t device, BfCudaShardRank **out) {
  if (!path || !out) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  *out = nullptr;
  BfCudaPlan *plan = nullptr;
  BfCudaStatus st = bfCudaPlanLoad(path, device, &plan);
  return st == BF_CUDA_OK ? wrapRankPlan(plan, out) : st;
}

extern "C" BfCudaPlan *bfCudaShardRankGetPlan(BfCudaShardRank *R) { return R ? R->plan : nullptr; }
extern "C" int64_t bfCudaShardRankSentRows(const BfCudaShardRank *R) { return R ? R->sentRows : 0; }

extern "C" BfCudaStatus bfCudaShardRankGetHandle(BfCudaShardRank *R, BfCudaShardHandle *h) {
  if (!R || !h) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  std::memset(h, 0, sizeof *h);
  h->pid = (int64_t)getpid();
  h->device = R->device;
  h->rank = R->rank;
  h->wsBytes = (int64_t)R->wsBytes;
  h->wsPtr = R->ws;
  h->flagsPtr = R->flags;
  if (R->world == 1) return BF_CUDA_OK;
  DeviceGuard g(R->device);
  static_assert(sizeof(cudaIpcMemHandle_t) <= sizeof h->wsIpc, "IPC handle size");
  cudaIpcMemHandle_t a, b;
  // (IPC export can be unavailable — e.g. inside some containers; same-process peers do not need it)
  if (cudaIpcGetMemHandle(&a, R->ws) == cudaSuccess && cudaIpcGetMemHandle(&b, R->flags) == cudaSuccess) {
    std::memcpy(h->wsIpc, &a, sizeof a);
    std::memcpy(h->flagsIpc, &b, sizeof b);
    h->hasIpc = 1;
  } else {
    cudaGetLastError();
  }
  return BF_CUDA_OK;
}

// all[r] = the handle rank r exported, r = 0 .. world - 1 (all[rank] is this rank's own).
extern "C" BfCudaStatus bfCudaShardRankConnect(BfCudaShardRank *R, const BfCudaShardHandle *all) {
  if (!R || !all) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  if (R->world == 1) { R->connected = true; return BF_CUDA_OK; }
  if (R->connected) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "already connected");
  DeviceGuard g(R->device);
  const int64_t me = (int64_t)getpid();
  for (int r = 0; r < R->world; ++r) {
    const BfCudaShardHandle &h = all[r];
    if (h.rank != r) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "handles are not in rank order");
    if (h.wsBytes != (int64_t)R->wsBytes) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "ranks disagree on the workspace layout (different operators?)");
    if (r == R->rank) { R->host.peerW[r] = R->ws; R->host.peerFlags[r] = R->flags; continue; }
    if (h.pid == me) {
      if (h.device != R->device) { // same process, another GPU: plain peer access
        int can = 0;
        BF_CUDA_TRY(cudaDeviceCanAccessPeer(&can, R->device, h.device));
        if (!can) return fail(BF_CUDA_ERROR_RUNTIME, "GPUs " + std::to_string(R->device) + " and " + std::to_string(h.device) + " cannot access each other's memory");
        cudaError_t e = cudaDeviceEnablePeerAccess(h.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) BF_CUDA_TRY(e);
        cudaGetLastError();
      }
      R->host.peerW[r] = static_cast<unsigned char *>(h.wsPtr);
      R->host.peerFlags[r] = static_cast<uint32_t *>(h.flagsPtr);
    } else {
      if (!h.hasIpc) return fail(BF_CUDA_ERROR_RUNTIME, "peer lives in another process and exported no CUDA IPC handle");
      cudaIpcMemHandle_t a, b;
      std::memcpy(&a, h.wsIpc, sizeof a);
      std::memcpy(&b, h.flagsIpc, sizeof b);
      BF_CUDA_TRY(cudaIpcOpenMemHandle(&R->opened[r][0], a, cudaIpcMemLazyEnablePeerAccess));
      BF_CUDA_TRY(cudaIpcOpenMemHandle(&R->opened[r][1], b, cudaIpcMemLazyEnablePeerAccess));
      R->host.peerW[r] = static_cast<unsigned char *>(R->opened[r][0]);
      R->host.peerFlags[r] = static_cast<uint32_t *>(R->opened[r][1]);
    }
  }
  BF_CUDA_TRY(cudaMemcpy(R->dev, &R->host, sizeof(ShardLaunch), cudaMemcpyHostToDevice));
  BF_CUDA_TRY(cudaStreamSynchronize(cudaStreamLegacy)); // landed before a stage kernel on a non-blocking stream reads it
  R->plan->extWs = R->ws;
  R->plan->extWsBytes = R->wsBytes;
  R->plan->shardDev = R->dev;
  R->connected = true;
  return BF_CUDA_OK;
}

// Stages [stageBegin, stageEnd) of one apply (stageEnd < 0: all).  d_x: the FULL operand (numCols elements, dense,
// 16-byte aligned) on this rank's device; d_y: numRows elements, of which this rank writes the rows it owns.
// A call with stageBegin == 0 starts a new apply.  Asynchronous.
extern "C" BfCudaStatus bfCudaShardRankApplyStages(BfCudaShardRank *R, const void *d_x, void *d_y, int64_t stageBegin,
                                                   int64_t stageEnd, void *stream) {
  if (!R) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  if (R->world == 1) return bfCudaApplyDevice(R->plan, BF_CUDA_OP_N, d_x, 1, 1, d_y, 1, stream);
  if (!R->connected) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "bfCudaShardRankConnect has not been called");
  if (stageBegin == 0) ++R->plan->shardEpoch;
  if (stageEnd < 0) stageEnd = R->numStages;
  return bfCudaApplyStages(R->plan, BF_CUDA_OP_N, d_x, 1, 1, d_y, 1, stageBegin, stageEnd, stream);
}

extern "C" BfCudaStatus bfCudaShardRankApplyDevice(BfCudaShardRank *R, const void *d_x, void *d_y, void *stream) {
  return bfCudaShardRankApplyStages(R, d_x, d_y, 0, -1, stream);
}

// ---- all ranks in one process ---------------------------------------------------------------------------------
struct BfCudaComm {
  std::vector<int> devices;
};

extern "C" BfCudaStatus bfCudaCommInit(int32_t nGpus, const int32_t *devices, BfCudaComm **out) {
  if (!out || nGpus < 1 || nGpus > kMaxShardRanks) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "1 .. 16 GPUs");
  *out = nullptr;
  int nDev = 0;
  if (cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0) {
    cudaGetLastError();
    return fail(BF_CUDA_ERROR_NO_DEVICE, "no CUDA device visible; this library has no CPU fallback");
  }
  std::unique_ptr<BfCudaComm> c(new BfCudaComm());
  for (int i = 0; i < nGpus; ++i) {
    const int d = devices ? devices[i] : i;
    if (d < 0 || d >= nDev) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "device ordinal out of range (" + std::to_string(nDev) + " visible)");
    c->devices.push_back(d);
  }
  *out = c.release();
  return BF_CUDA_OK;
}
extern "C" BfCudaStatus bfCudaCommDestroy(BfCudaComm *c) { delete c; return BF_CUDA_OK; }
extern "C" int32_t bfCudaCommSize(const BfCudaComm *c) { return c ? (int32_t)c->devices.size() : 0; }

struct BfCudaShardedPlan {
  int mode = BF_CUDA_SHARD_BLOCK_ROWS, world = 1, dtype = BF_CUDA_REAL_F64;
  int64_t numRows = 0, numCols = 0, entries = 0;
  std::vector<int> devices;
  std::vector<int64_t> yCuts;
  std::vector<BfCudaShardRank *> ranks; // block rows
  std::vector<BfCudaPlan *> plans;      // RHS columns (replicas)
  std::vector<cudaStream_t> streams;    // one per rank
  std::vector<unsigned char *> dX, dY;  // host-API buffers per rank
  std::vector<size_t> dXBytes, dYBytes;
  bool sharedDevice = false;            // two ranks on one GPU: everything goes through ONE stream, in stage order
  unsigned char *pinX = nullptr, *pinY = nullptr; // pinned staging for pageable host operands
  size_t pinXBytes = 0, pinYBytes = 0;
};

extern "C" BfCudaStatus bfCudaShardedPlanDestroy(BfCudaShardedPlan *S) {
  if (!S) return BF_CUDA_OK;
  for (size_t r = 0; r < S->streams.size(); ++r) {
    DeviceGuard g(S->devices[r]);
    if (S->streams[r]) cudaStreamSynchronize(S->streams[r]);
  }
  for (BfCudaShardRank *R : S->ranks) bfCudaShardRankDestroy(R);
  for (BfCudaPlan *P : S->plans) bfCudaPlanDestroy(P);
  for (size_t r = 0; r < S->devices.size(); ++r) {
    DeviceGuard g(S->devices[r]);
    if (r < S->dX.size()) cudaFree(S->dX[r]);
    if (r < S->dY.size()) cudaFree(S->dY[r]);
    if (r < S->streams.size() && S->streams[r] && !(S->sharedDevice && r > 0)) cudaStreamDestroy(S->streams[r]);
  }
  if (S->pinX) cudaFreeHost(S->pinX);
  if (S->pinY) cudaFreeHost(S->pinY);
  delete S;
  return BF_CUDA_OK;
}

extern "C" BfCudaStatus bfCudaPlanShard(const BfCudaPlanDesc *desc, BfCudaComm *comm, int32_t mode, const double *fractions,
                                        BfCudaShardedPlan **out) {
  if (!desc || !comm || !out) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  if (mode != BF_CUDA_SHARD_RHS_COLS && mode != BF_CUDA_SHARD_BLOCK_ROWS) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "unknown sharding mode");
  *out = nullptr;
  std::unique_ptr<BfCudaShardedPlan> S(new BfCudaShardedPlan());
  S->mode = mode;
  S->world = (int)comm->devices.size();
  S->devices = comm->devices;
  S->dtype = desc->dtype;
  S->numRows = desc->numRows; S->numCols = desc->numCols;
  const size_t W = (size_t)S->world;
  for (size_t a = 0; a < W; ++a)
    for (size_t b = a + 1; b < W; ++b)
      if (S->devices[a] == S->devices[b]) S->sharedDevice = true;
  S->streams.assign(W, nullptr);
  S->dX.assign(W, nullptr); S->dY.assign(W, nullptr);
  S->dXBytes.assign(W, 0); S->dYBytes.assign(W, 0);
  auto bail = [&](BfCudaStatus st) { bfCudaShardedPlanDestroy(S.release()); return st; };
  for (size_t r = 0; r < W; ++r) {
    DeviceGuard g(S->devices[r]);
    if (S->sharedDevice && r > 0) { S->streams[r] = S->streams[0]; continue; }
    cudaError_t e = cudaStreamCreateWithFlags(&S->streams[r], cudaStreamNonBlocking);
    if (e != cudaSuccess) return bail(fail(BF_CUDA_ERROR_RUNTIME, std::string("stream: ") + cudaGetErrorString(e)));
  }
  if (mode == BF_CUDA_SHARD_RHS_COLS) {
    for (size_t r = 0; r < W; ++r) {
      BfCudaPlanDesc d = *desc;
      d.device = S->devices[r];
      d.shardWorld = 0; d.shardRank = 0;
      BfCudaPlan *P = nullptr;
      BfCudaStatus st = bfCudaPlanCreate(&d, &P);
      if (st != BF_CUDA_OK) return bail(st);
      S->plans.push_back(P);
    }
    S->entries = S->plans[0]->dir[0].host.entries;
    *out = S.release();
    return BF_CUDA_OK;
  }
  BfCudaShardDescs *sd = nullptr;
  BfCudaStatus st = bfCudaShardDescsCreate(desc, S->world, fractions, &sd);
  if (st != BF_CUDA_OK) return bail(st);
  S->yCuts = sd->S.yCuts;
  for (size_t r = 0; r < W && st == BF_CUDA_OK; ++r) {
    BfCudaPlanDesc d = *bfCudaShardDescsGet(sd, (int)r);
    d.device = S->devices[r];
    BfCudaShardRank *R = nullptr;
    st = bfCudaShardRankCreate(&d, &R);
    if (st == BF_CUDA_OK) S->ranks.push_back(R);
  }
  bfCudaShardDescsFree(sd);
  if (st != BF_CUDA_OK) return bail(st);
  std::vector<BfCudaShardHandle> handles(W);
  for (size_t r = 0; r < W; ++r) {
    st = bfCudaShardRankGetHandle(S->ranks[r], &handles[r]);
    if (st != BF_CUDA_OK) return bail(st);
  }
  for (size_t r = 0; r < W; ++r) {
    st = bfCudaShardRankConnect(S->ranks[r], handles.data());
    if (st != BF_CUDA_OK) return bail(st);
    S->entries += S->ranks[r]->plan->dir[0].host.entries;
  }
  *out = S.release();
  return BF_CUDA_OK;
}

extern "C" BfCudaStatus bfCudaShardedPlanGetInfo(const BfCudaShardedPlan *S, BfCudaShardedInfo *info) {
  if (!S || !info) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  std::memset(info, 0, sizeof *info);
  info->world = S->world; info->mode = S->mode;
  info->numRows = S->numRows; info->numCols = S->numCols;
  info->numEntries = S->entries;
  for (BfCudaShardRank *R : S->ranks) {
    const Direction &D = R->plan->dir[0].host;
    info->maxRankEntries = std::max(info->maxRankEntries, D.entries);
    info->poolBytesMax = std::max<int64_t>(info->poolBytesMax, (int64_t)D.pool16 * 16);
    info->sentRowsPerApply += R->sentRows;
    info->numStages = (int64_t)D.stages.size();
  }
  for (BfCudaPlan *P : S->plans) {
    const Direction &D = P->dir[0].host;
    info->maxRankEntries = D.entries;
    info->poolBytesMax = std::max<int64_t>(info->poolBytesMax, (int64_t)D.pool16 * 16);
    info->numStages = (int64_t)D.stages.size();
  }
  for (size_t i = 0; i < S->yCuts.size() && i < 17; ++i) info->yCuts[i] = S->yCuts[i];
  return BF_CUDA_OK;
}

extern "C" BfCudaPlan *bfCudaShardedPlanGetRankPlan(BfCudaShardedPlan *S, int32_t rank) {
  if (!S || rank < 0 || rank >= S->world) return nullptr;
  return S->mode == BF_CUDA_SHARD_RHS_COLS ? S->plans[(size_t)rank] : S->ranks[(size_t)rank]->plan;
}

extern "C" void *bfCudaShardedRankStream(BfCudaShardedPlan *S, int32_t rank) {
  return S && rank >= 0 && rank < S->world ? (void *)S->streams[(size_t)rank] : nullptr;
}

extern "C" BfCudaStatus bfCudaShardedSynchronize(BfCudaShardedPlan *S) {
  if (!S) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  for (size_t r = 0; r < S->streams.size(); ++r) {
    DeviceGuard g(S->devices[r]);
    BF_CUDA_TRY(cudaStreamSynchronize(S->streams[r]));
  }
  return BF_CUDA_OK;
}

// Stage-major over the ranks: stage s of every rank is enqueued before stage s + 1 of any, so that no device
// starts an apply more than a few launches after the first one (a rank-major order would delay rank r by r x
// numStages launches), and — where ranks share a GPU and therefore a stream — no kernel ever waits for one
// queued behind it.
extern "C" BfCudaStatus bfCudaShardedApplyDevice(BfCudaShardedPlan *S, const void *const *d_x, void *const *d_y) {
  if (!S || !d_x || !d_y) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null argument");
  if (S->mode != BF_CUDA_SHARD_BLOCK_ROWS) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "device-resident sharded apply is for block rows (RHS columns: apply the rank plans)");
  const int nS = S->world == 1 ? 1 : S->ranks[0]->numStages;
  for (int s = 0; s < nS; ++s)
    for (size_t r = 0; r < S->ranks.size(); ++r) {
      DeviceGuard g(S->devices[r]);
      BfCudaStatus st = S->world == 1 ? bfCudaShardRankApplyDevice(S->ranks[r], d_x[r], d_y[r], S->streams[r])
                                      : bfCudaShardRankApplyStages(S->ranks[r], d_x[r], d_y[r], s, s + 1, S->streams[r]);
      if (st != BF_CUDA_OK) return st;
    }
  return BF_CUDA_OK;
}

static BfCudaStatus growDevice(unsigned char **p, size_t *have, size_t want) {
  if (*have >= want) return BF_CUDA_OK;
  cudaFree(*p);
  *p = nullptr; *have = 0;
  BF_CUDA_TRY(cudaMalloc(p, want));
  *have = want;
  return BF_CUDA_OK;
}
static BfCudaStatus growPinned(unsigned char **p, size_t *have, size_t want) {
  if (*have >= want) return BF_CUDA_OK;
  if (*p) cudaFreeHost(*p);
  *p = nullptr; *have = 0;
  BF_CUDA_TRY(cudaHostAlloc(p, want, cudaHostAllocPortable));
  *have = want;
  return BF_CUDA_OK;
}

#include <thread>

extern "C" BfCudaStatus bfCudaShardedApplyHost(BfCudaShardedPlan *S, int32_t op, const void *x, int64_t ldx, int64_t nrhs,
                                               void *y, int64_t ldy) {
  if (!S) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null plan");
  if (nrhs < 0 || (nrhs > 0 && (ldx < nrhs || ldy < nrhs))) return fail(BF_CUDA_ERROR_INCOMPATIBLE_SHAPES, "ldx / ldy smaller than nrhs");
  if (nrhs == 0) return BF_CUDA_OK;
  if (!x || !y) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null operand");
  const size_t es = S->dtype == BF_CUDA_COMPLEX_F64 ? 16 : 8;
  const size_t W = (size_t)S->world;
  if (S->mode == BF_CUDA_SHARD_RHS_COLS) {
    // device r applies columns [c0, c1): one host thread per device (bfCudaApplyHost is synchronous), no data
    // crosses GPUs
    std::vector<BfCudaStatus> st(W, BF_CUDA_OK);
    std::vector<std::thread> th;
    for (size_t r = 0; r < W; ++r) {
      const int64_t c0 = nrhs * (int64_t)r / (int64_t)W, c1 = nrhs * (int64_t)(r + 1) / (int64_t)W;
      if (c1 == c0) continue;
      const unsigned char *xr = static_cast<const unsigned char *>(x) + (size_t)c0 * es;
      unsigned char *yr = static_cast<unsigned char *>(y) + (size_t)c0 * es;
      th.emplace_back([=, &st]() { st[r] = bfCudaApplyHost(S->plans[r], op, xr, ldx, c1 - c0, yr, ldy); });
    }
    for (std::thread &t : th) t.join();
    for (size_t r = 0; r < W; ++r)
      if (st[r] != BF_CUDA_OK) return fail(st[r], "RHS-column shard " + std::to_string(r) + " failed");
    return BF_CUDA_OK;
  }
  if (op != BF_CUDA_OP_N) return fail(BF_CUDA_ERROR_NOT_IMPLEMENTED, "block-row shards are forward only");
  const size_t nIn = (size_t)S->numCols, nOut = (size_t)S->numRows;
  BfCudaStatus st = growPinned(&S->pinX, &S->pinXBytes, std::max<size_t>(nIn * es, 16));
  if (st == BF_CUDA_OK) st = growPinned(&S->pinY, &S->pinYBytes, std::max<size_t>(nOut * es, 16));
  for (size_t r = 0; r < W && st == BF_CUDA_OK; ++r) {
    DeviceGuard g(S->devices[r]);
    st = growDevice(&S->dX[r], &S->dXBytes[r], std::max<size_t>(nIn * es, 16));
    if (st == BF_CUDA_OK) st = growDevice(&S->dY[r], &S->dYBytes[r], std::max<size_t>(nOut * es, 16));
  }
  if (st != BF_CUDA_OK) return st;
  std::vector<const void *> xs(W);
  std::vector<void *> ys(W);
  for (size_t r = 0; r < W; ++r) { xs[r] = S->dX[r]; ys[r] = S->dY[r]; }
  const unsigned char *xb = static_cast<const unsigned char *>(x);
  unsigned char *yb = static_cast<unsigned char *>(y);
  for (int64_t c = 0; c < nrhs; ++c) {
    // column c -> pinned -> every device (X is replicated: n x 1 is tiny next to the factors)
    if (ldx == 1) std::memcpy(S->pinX, xb, nIn * es);
    else for (size_t i = 0; i < nIn; ++i) std::memcpy(S->pinX + i * es, xb + ((size_t)i * (size_t)ldx + (size_t)c) * es, es);
    for (size_t r = 0; r < W; ++r) {
      DeviceGuard g(S->devices[r]);
      BF_CUDA_TRY(cudaMemcpyAsync(S->dX[r], S->pinX, nIn * es, cudaMemcpyHostToDevice, S->streams[r]));
    }
    st = bfCudaShardedApplyDevice(S, xs.data(), ys.data());
    if (st != BF_CUDA_OK) return st;
    for (size_t r = 0; r < W; ++r) { // every rank returns the rows it owns
      const size_t r0 = W == 1 ? 0 : (size_t)S->yCuts[r], r1 = W == 1 ? nOut : (size_t)S->yCuts[r + 1];
      if (r1 == r0) continue;
      DeviceGuard g(S->devices[r]);
      BF_CUDA_TRY(cudaMemcpyAsync(S->pinY + r0 * es, S->dY[r] + r0 * es, (r1 - r0) * es, cudaMemcpyDeviceToHost, S->streams[r]));
    }
    st = bfCudaShardedSynchronize(S);
    if (st != BF_CUDA_OK) return st;
    if (ldy == 1) std::memcpy(yb, S->pinY, nOut * es);
    else for (size_t i = 0; i < nOut; ++i) std::memcpy(yb + ((size_t)i * (size_t)ldy + (size_t)c) * es, S->pinY + i * es, es);
  }
  return BF_CUDA_OK;
}
