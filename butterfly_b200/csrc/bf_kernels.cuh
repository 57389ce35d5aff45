// bf_kernels.cuh — launch interface between the C ABI (bf_capi.cu) and the sm_100a kernels.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "bf_plan.hpp"

namespace bfc {

// Everything one stage launch needs; passed by value as a __grid_constant__ kernel parameter.
struct StageLaunch {
  const DevItem *items;
  const DevTile *tiles;
  const uint32_t *qItemBegin; // [numQueues + 1]
  const uint32_t *qTileBegin; // [numQueues + 1]
  const uint64_t *qPool16;    // [numQueues + 1]
  const unsigned char *pool;  // byte 0 of the value pool
  unsigned char *base[3];     // X, Y, workspace
  long long ld[3];            // row strides in ELEMENTS
  const int32_t *outPerm;     // NULL, or: row r of the result is stored at row outPerm[r]
  int numQueues;
  int nrhs;
  int colBegin;               // first RHS column this launch handles (wide kernel)
  int pdl;                    // 1-RHS kernel: launched with programmatic stream serialization (small operators)
  // fused single-launch apply (FusedTables): dependency words parallel to items / tiles, the counter set of
  // this apply and the one to zero for the next
  const ItemDep *itemDeps;
  const TileDep *tileDeps;
  uint32_t *cnt, *cntNext;
  uint32_t numCounters;
  // block-row shard (one rank of a sharded operator; bf_comm.cu fills these, see ShardLaunch)
  const struct ShardLaunch *shard; // device copy, NULL for unsharded plans
  const uint32_t *itemPeers;       // per item of this stage: ranks that read its rows
  int stage;
  uint32_t epoch;                  // apply counter, >= 1
};

// What a stage kernel of a block-row shard needs to hand rows to its peers: it stores the rows another rank reads
// straight into that rank's workspace (peer memory over NVLink), and the LAST warp of the grid to finish posts one
// flag per peer — flags[stage][this rank] = epoch in the peer's flag array; the peer's next stage spins on it.  No
// collective, no host round trip: all stages of all ranks are enqueued at once.
constexpr int kMaxRanks = 16;
struct ShardLaunch {
  unsigned char *peerW[kMaxRanks];  // workspace base of every rank (entry `rank`: my own)
  uint32_t *peerFlags[kMaxRanks];   // flag array of every rank: [numStages][world]
  uint32_t *done;                   // [numStages] warps of my grid that finished stage s (monotonic over applies)
  uint32_t recvMask[32];            // per stage: ranks that push rows to me during that stage
  uint32_t sendMask[32];            // per stage: ranks I push rows to (they wait for my flag of that stage)
  int world, rank, numStages;
  uint32_t totalWarps;              // warps per stage grid
};

// Shared-memory geometry of the 1-RHS kernel (one ring of kGemvSlots slots per warp).
constexpr int kGemvSlots = 2;
constexpr int kGemvXBytes = 1056;                        // x segment: 64 c128 / 128 f64 (+16 B slack)
constexpr int kGemvSlotBytes = kTileBytes + kGemvXBytes; // 9248
constexpr int kGemvSmemBytes = kWarpsPerCta * kGemvSlots * (kGemvSlotBytes + 8) + 128;

// Wide (many-RHS) kernel: CTA-wide ring of A tiles; one CTA owns kWideQueuesPerCta consecutive queues.
constexpr int kWideSlots = 4;
constexpr int kWideThreads = 128;       // 4 warps: warp = row group, lane = column group
constexpr int kWideQueuesPerCta = 2;
constexpr int kWideSmemBytes = kWideSlots * (kTileBytes + 8) + 128;
// DMMA (complex many-RHS) kernel: one byte ring per CTA holding A tiles + operand row segments.
constexpr int kMmaThreads = 160;                  // 4 consumer warps + 1 producer warp
constexpr int kMmaBars = 8;                       // tiles in flight
constexpr int kMmaRingBytes = 96 * 1024;          // 2 CTAs / SM
constexpr int kMmaSmemBytes = kMmaRingBytes + kMmaBars * (8 + 8 + 16) + 64;

// 1 RHS: y = A x, one warp per queue, TMA-staged tiles, shuffle reductions.
cudaError_t launchGemvStage(int dtype, const StageLaunch &a, int numSMs, cudaStream_t stream);
// 1 RHS, one stage of a block-row shard (a.shard != NULL): peer stores + flags.
cudaError_t launchGemvShardStage(int dtype, const StageLaunch &a, cudaStream_t stream);
// 1 RHS, all stages in one cooperative launch (`a` carries the fused tables and the counters).
cudaError_t launchGemvFused(int dtype, const StageLaunch &a, cudaStream_t stream);
// nrhs >= 2: one CTA per kWarpsPerCta queues, columns across threads.
cudaError_t launchWideStage(int dtype, const StageLaunch &a, int numSMs, cudaStream_t stream);
cudaError_t configureKernels(); // opt in to the large dynamic shared memory carve-outs

// dst[(perm ? perm[i] : i) * ldDst + c] = src[i * ldSrc + c]  (elements of elemDoubles doubles).
// Fuses the caller-side bfMatPermuteRows (src/mat_dense_complex.c:741-774) into the staging copy.
cudaError_t launchPermuteRows(const double *src, long long ldSrc, double *dst, long long ldDst,
                              const int32_t *perm, long long rows, int nrhs, int elemDoubles,
                              cudaStream_t stream);

// Synthetic pool fill (bfCudaPlanFillSynthetic): one entry per tile.
struct FillPiece {
  uint64_t pool16;    // where the tile lives
  int64_t rawIndex;   // raw op
  int32_t row0, col0; // piece origin in the direction's acting matrix
  int32_t storedCols;
  uint16_t rows, kc;
  uint8_t swapRC, conj, kind, pad;
  double scale;
};
cudaError_t launchFillSynthetic(int dtype, unsigned char *pool, const FillPiece *pieces, long long numPieces,
                                uint64_t seed, cudaStream_t stream);

} // namespace bfc
