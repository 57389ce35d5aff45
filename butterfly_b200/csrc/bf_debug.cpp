// bf_debug.cpp — host-only access to the compiled tables (include/bf_cuda_debug.h).
#include "bf_cuda_debug.h"

#include <cstring>
#include <string>
#include <vector>

#include "bf_plan.hpp"

using namespace bfc;

namespace {
struct Owner {
  Direction D;
  std::vector<int64_t> stageItemBegin, stageTileBegin, tileOp;
  std::vector<DevItem> items;
  std::vector<DevTile> tiles;
  std::vector<uint64_t> tilePool16, qPool16;
  std::vector<uint32_t> qItemBegin, qTileBegin;
  std::vector<int32_t> tileRow0, tileCol0;
  std::vector<unsigned char> pool;
  std::vector<uint32_t> itemPeers, recvMask, sendMask;
  BfCudaDebugTables t;
};
} // namespace

extern "C" BfCudaStatus bfCudaDebugCompile(const BfCudaPlanDesc *desc, int32_t op, int32_t numQueues,
                                           BfCudaDebugTables **out) {
  if (!desc || !out || numQueues <= 0) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  *out = nullptr;
  Owner *o = new Owner();
  std::string err;
  BfCudaStatus st = compileDirection(*desc, op, numQueues, o->D, err);
  if (st != BF_CUDA_OK) { delete o; return st; }
  const Direction &D = o->D;
  for (const Stage &S : D.stages) {
    o->stageItemBegin.push_back((int64_t)o->items.size());
    o->stageTileBegin.push_back((int64_t)o->tiles.size());
    o->items.insert(o->items.end(), S.items.begin(), S.items.end());
    o->tiles.insert(o->tiles.end(), S.tiles.begin(), S.tiles.end());
    o->tilePool16.insert(o->tilePool16.end(), S.tilePool16.begin(), S.tilePool16.end());
    o->qItemBegin.insert(o->qItemBegin.end(), S.qItemBegin.begin(), S.qItemBegin.end());
    o->qTileBegin.insert(o->qTileBegin.end(), S.qTileBegin.begin(), S.qTileBegin.end());
    o->qPool16.insert(o->qPool16.end(), S.qPool16.begin(), S.qPool16.end());
    for (const PieceRef &p : S.pieces) {
      o->tileOp.push_back(D.ops[(size_t)p.op].rawIndex);
      o->tileRow0.push_back(p.row0);
      o->tileCol0.push_back(p.col0);
    }
  }
  o->stageItemBegin.push_back((int64_t)o->items.size());
  o->stageTileBegin.push_back((int64_t)o->tiles.size());
  bool anyVal = false; // value-less (structure-only) operators get tables without a packed pool
  for (const DirOp &d : D.ops) if (d.val) { anyVal = true; break; }
  if (anyVal) {
    o->pool.assign((size_t)D.pool16 * 16 + 16, 0);
    packPool(D, 0, D.stages.size(), o->pool.data());
  }
  BfCudaDebugTables &t = o->t;
  std::memset(&t, 0, sizeof t);
  t.dtype = D.dtype; t.op = op;
  t.numStages = (int32_t)D.stages.size(); t.numQueues = numQueues;
  t.rowsIn = D.rowsIn; t.rowsOut = D.rowsOut; t.wsRows = D.wsRows;
  t.numItems = (int64_t)o->items.size(); t.numTiles = (int64_t)o->tiles.size();
  t.stageItemBegin = o->stageItemBegin.data(); t.stageTileBegin = o->stageTileBegin.data();
  t.items = o->items.data(); t.tiles = o->tiles.data();
  t.tilePool16 = o->tilePool16.data();
  t.qItemBegin = o->qItemBegin.data(); t.qTileBegin = o->qTileBegin.data(); t.qPool16 = o->qPool16.data();
  t.tileOp = o->tileOp.data(); t.tileRow0 = o->tileRow0.data(); t.tileCol0 = o->tileCol0.data();
  t.pool16 = D.pool16; t.pool = anyVal ? o->pool.data() : nullptr;
  t.wsArena0Rows = D.wsArena[0]; t.wsArena1Rows = D.wsArena[1];
  t.owner = o;
  const FusedTables &F = D.fused;
  t.fusedEnabled = F.enabled ? 1 : 0;
  t.fusedNumCounters = F.numCounters;
  t.fusedWsRows = F.wsRows;
  t.fusedNumItems = (int64_t)F.items.size(); t.fusedNumTiles = (int64_t)F.tiles.size();
  t.fusedItems = F.items.data(); t.fusedItemDeps = F.itemDeps.data();
  t.fusedTiles = F.tiles.data(); t.fusedTileDeps = F.tileDeps.data();
  t.fusedQItemBegin = F.qItemBegin.data(); t.fusedQTileBegin = F.qTileBegin.data();
  if (D.shardWorld > 1) {
    for (const Stage &S : D.stages) {
      o->itemPeers.insert(o->itemPeers.end(), S.itemPeers.begin(), S.itemPeers.end());
      o->recvMask.push_back(S.recvMask);
      o->sendMask.push_back(S.sendMask);
    }
    t.itemPeers = o->itemPeers.data();
    t.stageRecvMask = o->recvMask.data();
    t.stageSendMask = o->sendMask.data();
  }
  *out = &t;
  return BF_CUDA_OK;
}

extern "C" void bfCudaDebugFree(BfCudaDebugTables *t) {
  if (t) delete static_cast<Owner *>(t->owner);
}
