// bf_plan.hpp — host-side plan compiler for the butterfly apply (no CUDA in this header).
//
// The raw-op list of include/bf_cuda.h (one leaf of the reference BfMat tree per op, in the
// reference's own traversal order) is compiled into a STAGE-MAJOR packed layout:
//
//   stage  = all leaves whose operand is available after the same number of dependent steps
//            (stage 0: near-field dense leaves + first butterfly factors; stage s: the s-th applied
//            factor of every MatProduct, src/mat_product.c:225-238, batched together)
//   item   = a run of <= R output rows of one virtual buffer that a single warp (1 RHS) or CTA
//            (many RHS) produces with NO atomics: it sums all of its tiles, in reference order
//            (src/mat_block_coo.c:404-418), and stores once
//   tile   = an (item rows) x (<= KC columns) piece of one leaf, re-packed contiguously into the
//            value pool in exactly the order the kernels consume it, so that every tile is ONE
//            bulk (TMA) copy and every warp streams a contiguous span of HBM
//
// DESIGN.md §3 describes the layout; tests/test_plan_tables.py interprets these tables on the
// CPU to pin the index handling bit-exactly without a GPU.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "bf_cuda.h"

namespace bfc {

// ---- tiling constants shared by the compiler and the kernels --------------------------------
constexpr int kWarpsPerCta = 12;      // gemv kernel: one work queue per warp
constexpr int kTileBytes = 8192;      // upper bound of one packed tile
constexpr int kMaxRowsComplex = 32;   // R for complex items (8 lanes per row, 8 passes)
constexpr int kMaxRowsReal = 16;      // R for real items (16 lanes per row, 8 passes)
constexpr int kMaxKcComplex = 32;     // KC upper bound (bounds the operand rows the DMMA kernel stages per tile)
constexpr int kMaxKcReal = 128;       // (8 register chunks of 16 lanes)

// Address of a row in one of three spaces, packed in 32 bits: space << 30 | row.
enum : uint32_t { kSpaceX = 0, kSpaceY = 1, kSpaceW = 2 };
constexpr uint32_t kRowMask = (1u << 30) - 1;
inline uint32_t packAddr(uint32_t space, int64_t row) { return (space << 30) | (uint32_t)row; }

// Special tile kinds are encoded in DevTile::kc.
constexpr uint16_t kKcIdentity = 0xFFFF; // y[r] += x[r]           (no pool bytes)
constexpr uint16_t kKcDiag = 0xFFFE;     // y[r] += d[r] * x[r]    (pool: rows doubles, padded to 16 B)

constexpr uint16_t kKcJump = 0xFFFD;     // fused tables only: the queue's pool cursor moves to 16 * (rows << 32 | inAddr)

constexpr uint16_t kItemAccumulate = 1;  // rows already hold a partial result from an earlier stage
constexpr uint16_t kItemJump = 2;        // fused tables only: pseudo item (rows == 0) whose single tile is a kKcJump
constexpr uint32_t kNoCounter = 0xFFFFFFFFu;

struct DevItem {      // 16 B
  uint32_t outAddr;   // first output row
  uint32_t tileBegin; // index into the stage's tile array
  uint16_t rows;      // 1..R
  uint16_t flags;     // kItemAccumulate
  uint32_t numTiles;  // 0 = store zeros (rows no leaf covers, src/mat_block_diag.c:384)
};
struct DevTile {      // 8 B
  uint32_t inAddr;    // first operand row
  uint16_t kc;        // columns of this piece, or kKcIdentity / kKcDiag
  uint16_t rows;      // == owning item's rows (lets the prefetcher run ahead of the consumer)
};
static_assert(sizeof(DevItem) == 16 && sizeof(DevTile) == 8, "table ABI");

// ---- fused (single-launch) tables ---------------------------------------------------------------
// The 1-RHS apply can run as ONE persistent kernel: warp q walks the concatenation of its queue over all
// stages, and the stage barrier is replaced by readiness counters.  The tables are a queue-major copy of
// the stage-major ones with
//   * workspace rows in a FLAT layout (every intermediate buffer has rows of its own: nothing is reused
//     inside an apply, so there are no write-after-read hazards and only true dependencies remain; at one
//     RHS column the extra rows cost megabytes),
//   * one counter per written workspace buffer and per result segment that several stages accumulate into;
//     an item bumps its counter after its store, a tile waits until every writer of its operand buffer has
//     done so, an accumulating item until the earlier stages' writers have,
//   * a jump pseudo item / tile wherever the queue's pool span is not contiguous (stage boundaries).
struct ItemDep {        // 8 B, parallel to the fused items
  uint32_t sig;         // counter bumped after the store (kNoCounter: nobody waits for these rows)
  uint32_t waitTarget;  // accumulating items: value the same counter must have reached before the read-modify-write
};
struct TileDep {        // 8 B, parallel to the fused tiles
  uint32_t idx;         // counter of the operand's buffer (kNoCounter: the operand is X)
  uint32_t total;       // its final value = number of items that write that buffer
};
struct FusedTables {
  bool enabled = false;
  std::vector<DevItem> items;       // tileBegin indexes `tiles` below
  std::vector<ItemDep> itemDeps;
  std::vector<DevTile> tiles;
  std::vector<TileDep> tileDeps;
  std::vector<uint32_t> qItemBegin; // [numQueues + 1]
  std::vector<uint32_t> qTileBegin; // [numQueues + 1]
  uint32_t numCounters = 0;
  int64_t wsRows = 0;               // rows of the flat workspace
};

// Where a tile's values come from (host only; used by packing and by the synthetic fill).
struct PieceRef {
  int64_t op;    // index into Direction::ops
  int32_t row0;  // first row of the piece inside the leaf's ACTING matrix for this direction
  int32_t col0;  // first column
};

// One leaf as it acts in a given direction: out[outBuf][outOff:+m] += D * in[inBuf][inOff:+k].
struct DirOp {
  int32_t kind;      // BfCudaLeafKind
  uint8_t swapRC;    // D(p,q) = stored(q,p) instead of stored(p,q)
  uint8_t conj;      // D takes the complex conjugate of the stored value
  uint8_t remote;    // BF_CUDA_LEAF_REMOTE: part of the structure, applied by another rank
  uint8_t owner;     // sharded descriptors: the rank that applies the leaf
  int64_t rawIndex;  // index in the BfCudaPlanDesc op list (synthetic values hash on this)
  int64_t inBuf, inOff, outBuf, outOff, m, k;
  int64_t ld;        // stored row stride (elements)
  int64_t storedCols;
  const void *val;
  int32_t stage;
};

struct Stage {
  std::vector<DevItem> items;
  std::vector<DevTile> tiles;
  std::vector<PieceRef> pieces;       // one per tile
  std::vector<uint32_t> qItemBegin;   // [numQueues + 1]
  std::vector<uint32_t> qTileBegin;   // [numQueues + 1]
  std::vector<uint64_t> qPool16;      // [numQueues + 1] pool offset (16-byte units) where a queue starts
  std::vector<uint64_t> tilePool16;   // [numTiles] pool offset of every tile (host only)
  std::vector<int64_t> itemBuf, itemElem; // per item: buffer written, elementary interval (host only; fused tables)
  std::vector<int64_t> itemRow;       // per item: first row inside that buffer
  std::vector<int64_t> tileBuf, tileRow; // per tile: buffer read and first row read (-1 for jump tiles)
  // block-row shards (desc.shardWorld > 1): bit r of itemPeers[i] = rank r reads rows of item i, so the stage
  // kernel stores them into r's workspace too; sendMask / recvMask = ranks this stage pushes rows to / is
  // pushed rows by; sentRows = workspace rows pushed (a row pushed to two ranks counts twice)
  std::vector<uint32_t> itemPeers;
  uint32_t sendMask = 0, recvMask = 0;
  int64_t sentRows = 0;
  uint64_t poolBegin16 = 0, poolEnd16 = 0;
  int64_t entries = 0;                // dense entries applied in this stage
};

struct Direction {
  int32_t op = BF_CUDA_OP_N;
  int32_t dtype = BF_CUDA_REAL_F64;
  int64_t rowsIn = 0, rowsOut = 0;
  int64_t inBufId = 0, outBufId = 1;
  int numQueues = 0;
  std::vector<DirOp> ops;
  std::vector<Stage> stages;
  std::vector<int64_t> wsOffset;      // per virtual buffer: first row in the workspace (-1: X / Y)
  int64_t wsRows = 0;                 // workspace rows per RHS column
  int64_t wsArena[2] = {0, 0};        // rows of the two ping-pong arenas (the rest is dedicated)
  uint64_t pool16 = 0;                // total pool size, 16-byte units
  int64_t entries = 0;
  int64_t numItems = 0, numTiles = 0, maxTileK = 0, maxItemRows = 0;
  FusedTables fused;
  std::vector<std::vector<double>> ownedValues; // values of merged leaves (bf_plan.cpp step 0); dropped once the pool is packed
  int shardWorld = 0, shardRank = 0;  // > 1: one rank's share; the workspace is then flat (no row is reused inside
                                      // an apply: peers push into it at their own pace)
};

// Tiling helpers (the kernels static_assert against the same numbers).
inline int elemBytes(int dtype) { return dtype == BF_CUDA_COMPLEX_F64 ? 16 : 8; }
inline int maxItemRows(int dtype) { return dtype == BF_CUDA_COMPLEX_F64 ? kMaxRowsComplex : kMaxRowsReal; }
int kcForRows(int dtype, int rows);   // widest tile (columns) allowed for an item of `rows` rows
inline uint64_t tileBytes16(int dtype, int rows, int kc) { // packed size in 16-byte units
  if (kc == kKcIdentity) return 0;
  if (kc == kKcDiag) return ((uint64_t)rows * 8 + 15) / 16;
  return ((uint64_t)rows * kc * elemBytes(dtype) + 15) / 16;
}

// Compile one direction.  Returns BF_CUDA_OK or an error with `err` filled.
BfCudaStatus compileDirection(const BfCudaPlanDesc &desc, int32_t op, int numQueues, Direction &out,
                              std::string &err);

// Pack the values of stages [s0, s1) into dst (dst[0] corresponds to pool offset
// stages[s0].poolBegin16).  Ops with val == NULL are left zero.
void packPool(const Direction &dir, size_t s0, size_t s1, void *dst);
// The same for tiles [t0, t1) of one stage; dst[0] corresponds to pool offset base16 (<= tilePool16[t0]).  Ranges of
// different tiles write disjoint bytes, so several threads may pack one buffer.
void packTileRange(const Direction &dir, size_t stage, size_t t0, size_t t1, uint64_t base16, void *dst);

// Deterministic synthetic value of stored element `linear` (= storedRow * storedCols + storedCol)
// of raw op `rawIndex`; part 0 = real, 1 = imaginary; uniform in [-scale, scale).  Integer hash +
// one exact multiply, so host and device produce identical bits.  scale = syntheticScale(k).
#if defined(__CUDACC__)
__host__ __device__
#endif
inline double syntheticValue(uint64_t seed, int64_t rawIndex, int64_t linear, int32_t part, double scale) {
  uint64_t z = seed ^ (0x9E3779B97F4A7C15ull * (uint64_t)(rawIndex + 1));
  z += 0xBF58476D1CE4E5B9ull * (uint64_t)(2 * linear + part + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  double u = (double)(int64_t)(z >> 11) * (1.0 / 4503599627370496.0) - 1.0; // [-1, 1)
  return u * scale;
}
// 2^-round(log2(k)/2): a power of two near 1/sqrt(k), so that a k-term dot product of O(1)
// operands stays O(1) through many levels and the scaling itself is exact.
inline double syntheticScale(int64_t k) {
  int e = 0;
  while ((int64_t(1) << (e + 1)) <= k && e < 60) ++e; // floor(log2 k)
  return 1.0 / (double)(int64_t(1) << ((e + 1) / 2));
}

} // namespace bfc
