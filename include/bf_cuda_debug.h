/* bf_cuda_debug.h — host-only window into the plan compiler (no CUDA device needed).
 *
 * bfCudaDebugCompile runs exactly the compiler bfCudaPlanCreate runs and hands back the tables
 * and the packed value pool it would upload.  It exists so that the integer side of the plan
 * (stage assignment, item / tile row ranges, workspace offsets, queue partition, pool packing,
 * transposition and conjugation) can be pinned bit-exactly on a machine without a GPU:
 * tests/test_plan_tables.py interprets these tables with numpy and compares with the oracle.
 * Nothing on the product path calls it.
 */
#ifndef BF_CUDA_DEBUG_H
#define BF_CUDA_DEBUG_H
#include "bf_cuda.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct BfCudaDebugTables {
  int32_t dtype, op;
  int32_t numStages, numQueues;
  int64_t rowsIn, rowsOut, wsRows;
  int64_t numItems, numTiles;
  const int64_t *stageItemBegin; /* [numStages + 1] into items */
  const int64_t *stageTileBegin; /* [numStages + 1] into tiles (item.tileBegin is stage-relative) */
  const void *items;             /* numItems x {u32 outAddr, u32 tileBegin, u16 rows, u16 flags, u32 numTiles} */
  const void *tiles;             /* numTiles x {u32 inAddr, u16 kc, u16 rows} */
  const uint64_t *tilePool16;    /* [numTiles] offset of the tile's values in the pool, 16-byte units */
  const uint32_t *qItemBegin;    /* [numStages x (numQueues + 1)] stage-relative */
  const uint32_t *qTileBegin;    /* [numStages x (numQueues + 1)] */
  const uint64_t *qPool16;       /* [numStages x (numQueues + 1)] */
  const int64_t *tileOp;         /* [numTiles] raw op index the tile was cut from */
  const int32_t *tileRow0;       /* [numTiles] origin of the tile in the acting matrix */
  const int32_t *tileCol0;
  uint64_t pool16;               /* pool size in 16-byte units */
  const void *pool;              /* packed values (zeros where val == NULL); NULL if no op carries values */
  int64_t wsArena0Rows, wsArena1Rows; /* ping-pong arenas; rows past them are dedicated (bf_cuda.h) */
  void *owner;                   /* opaque */
  /* fused single-launch tables (bf_plan.hpp FusedTables): queue-major over all stages, flat workspace rows,
   * readiness counters.  fusedEnabled == 0: the operator has none (block-row shard, or too large). */
  int32_t fusedEnabled;
  uint32_t fusedNumCounters;
  int64_t fusedWsRows;
  int64_t fusedNumItems, fusedNumTiles;
  const void *fusedItems;        /* as `items`; tileBegin indexes fusedTiles; flags bit 1 = jump pseudo item */
  const void *fusedItemDeps;     /* fusedNumItems x {u32 sig, u32 waitTarget} */
  const void *fusedTiles;        /* as `tiles`; kc == 0xFFFD: pool cursor jumps to 16 * (rows << 32 | inAddr) */
  const void *fusedTileDeps;     /* fusedNumTiles x {u32 counter, u32 total} */
  const uint32_t *fusedQItemBegin; /* [numQueues + 1] */
  const uint32_t *fusedQTileBegin; /* [numQueues + 1] */
  /* block-row shards (desc->shardWorld > 1), NULL otherwise: per item (parallel to `items`) the ranks that read
   * its rows — the stage kernel stores them into those ranks' workspaces too — and per stage the ranks that push
   * rows to this one / that this one pushes rows to */
  const uint32_t *itemPeers;
  const uint32_t *stageRecvMask; /* [numStages] */
  const uint32_t *stageSendMask; /* [numStages] */
} BfCudaDebugTables;

BfCudaStatus bfCudaDebugCompile(const BfCudaPlanDesc *desc, int32_t op, int32_t numQueues,
                                BfCudaDebugTables **tables);
void bfCudaDebugFree(BfCudaDebugTables *tables);

#ifdef __cplusplus
}
#endif
#endif
