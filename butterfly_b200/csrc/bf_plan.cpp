// bf_plan.cpp — plan compiler: raw ops -> stage-major items / tiles / packed pool.  Host only.
#include "bf_plan.hpp"

#include <algorithm>
#include <array>
#include <climits>
#include <complex>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <map>
#include <queue>

namespace bfc {

int kcForRows(int dtype, int rows) {
  const bool cplx = dtype == BF_CUDA_COMPLEX_F64;
  const int lanesPerRow = cplx ? 8 : 16;
  const int cap = cplx ? kMaxKcComplex : kMaxKcReal;
  int kc = kTileBytes / (rows * elemBytes(dtype));
  kc -= kc % lanesPerRow;
  return std::max(lanesPerRow, std::min(cap, kc));
}

static bool leafStoredTransposed(const BfCudaRawOp &o, bool cplx) {
  // getCblasTranspose (src/mat_dense_complex.c:27-35): either bit => conjugate transpose;
  // real leaves: TRANS => transpose (src/mat_dense_real.c:1345-1358).
  if (cplx) return (o.flags & (BF_CUDA_LEAF_TRANS | BF_CUDA_LEAF_CONJ)) != 0;
  return (o.flags & BF_CUDA_LEAF_TRANS) != 0;
}

static bool sortItemsEnabled() { static const int v = std::getenv("BF_PLAN_SORT") ? std::atoi(std::getenv("BF_PLAN_SORT")) : 1; return v != 0; }

static bool deferHeavyEnabled() { const char *e = std::getenv("BF_PLAN_DEFER"); return !e || std::atoi(e) != 0; } // read per compile (A/B in one process)

// Cost model of the queue assignment, in byte equivalents: what a warp spends on a tile besides streaming its bytes.
// A tile is a serial chain for its warp (descriptor shuffles, mbarrier wait, shared-memory loads, a dependent FMA
// chain, the refill's bulk-copy issue): about as long as streaming 12 KiB once the stage is short enough that warps
// run alone at its end.  Measured (tools/plan_knob_ab.py, profiles/r2z_plan_tile_cost.txt): with 384 the queues of
// stage 0 at N = 16 384 carried 24 .. 92 tiles for near-equal bytes and the apply took 0.131 ms; with 12 288 they
// carry 38 .. 51 and it takes 0.107 ms (N = 65 536: 0.546 -> 0.527 ms; N = 524 288, HBM-bound: unchanged).
// Development knobs, read per compile: BF_PLAN_TILE_COST, BF_PLAN_ITEM_COST.
static uint64_t tileFixedCost() { const char *e = std::getenv("BF_PLAN_TILE_COST"); return e ? (uint64_t)std::atoll(e) : 12288ull; }

static uint64_t itemFixedCost() { const char *e = std::getenv("BF_PLAN_ITEM_COST"); return e ? (uint64_t)std::atoll(e) : 256ull; }

static uint64_t itemCost(int dtype, const Stage &S, const DevItem &it, uint64_t tileCost, uint64_t itemFixed) {
  uint64_t c = itemFixed + 32ull * it.rows;
  for (uint32_t t = it.tileBegin; t < it.tileBegin + it.numTiles; ++t) c += 16 * tileBytes16(dtype, S.tiles[t].rows, S.tiles[t].kc) + tileCost;
  return c;
}

// The items of a stage are independent (disjoint output rows), so their order — and which queue takes which — is
// free.  LPT: largest item first, each to the queue that is the lightest at that moment; the tables are then
// emitted queue by queue (inside a queue in the original order).  The slowest queue ends up at about
// max(mean, largest item), where a contiguous cut of the original order left it at 1.2 - 1.35x the mean on an 8-way
// row shard.  Tiles are re-ordered with their items (a queue walks the tile array and the pool sequentially).
// Returns qEnd[q] = one past the last item of queue q.
static std::vector<uint32_t> assignItemsLpt(int dtype, Stage &S, int numQueues) {
  const size_t nI = S.items.size();
  std::vector<uint32_t> qEnd((size_t)numQueues, (uint32_t)nI);
  if (nI == 0) return qEnd;
  std::vector<uint64_t> cost(nI);
  const uint64_t tileCost = tileFixedCost();
  const uint64_t itemFixed = itemFixedCost();
  for (size_t i = 0; i < nI; ++i) cost[i] = itemCost(dtype, S, S.items[i], tileCost, itemFixed);
  std::vector<uint32_t> order(nI);
  std::iota(order.begin(), order.end(), 0u);
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return cost[a] > cost[b]; });
  using Bin = std::pair<uint64_t, int>; // (load, queue): min-heap
  std::priority_queue<Bin, std::vector<Bin>, std::greater<Bin>> heap;
  for (int q = 0; q < numQueues; ++q) heap.push(Bin(0, q));
  std::vector<int> binOf(nI);
  for (uint32_t i : order) {
    Bin b = heap.top();
    heap.pop();
    binOf[i] = b.second;
    b.first += cost[i];
    heap.push(b);
  }
  std::vector<uint32_t> count((size_t)numQueues + 1, 0);
  for (size_t i = 0; i < nI; ++i) ++count[(size_t)binOf[i] + 1];
  for (int q = 0; q < numQueues; ++q) count[(size_t)q + 1] += count[(size_t)q];
  std::vector<uint32_t> slot(count.begin(), count.end() - 1), perm(nI);
  for (size_t i = 0; i < nI; ++i) perm[slot[(size_t)binOf[i]]++] = (uint32_t)i; // original order inside a queue
  for (int q = 0; q < numQueues; ++q) qEnd[(size_t)q] = count[(size_t)q + 1];
  Stage N;
  N.items.reserve(nI); N.itemBuf.reserve(nI); N.itemElem.reserve(nI); N.itemRow.reserve(nI);
  N.tiles.reserve(S.tiles.size()); N.pieces.reserve(S.tiles.size()); N.tileBuf.reserve(S.tiles.size()); N.tileRow.reserve(S.tiles.size());
  for (uint32_t i : perm) {
    DevItem it = S.items[i];
    const uint32_t tb = it.tileBegin;
    it.tileBegin = (uint32_t)N.tiles.size();
    for (uint32_t t = tb; t < tb + it.numTiles; ++t) {
      N.tiles.push_back(S.tiles[t]); N.pieces.push_back(S.pieces[t]);
      N.tileBuf.push_back(S.tileBuf[t]); N.tileRow.push_back(S.tileRow[t]);
    }
    N.items.push_back(it);
    N.itemBuf.push_back(S.itemBuf[i]); N.itemElem.push_back(S.itemElem[i]); N.itemRow.push_back(S.itemRow[i]);
  }
  S.items.swap(N.items); S.itemBuf.swap(N.itemBuf); S.itemElem.swap(N.itemElem); S.itemRow.swap(N.itemRow);
  S.tiles.swap(N.tiles); S.pieces.swap(N.pieces); S.tileBuf.swap(N.tileBuf); S.tileRow.swap(N.tileRow);
  return qEnd;
}

namespace {
struct Frag {       // one (leaf, elementary output interval) incidence
  int32_t stage;
  int64_t elem;     // global elementary-interval id (ordered by buffer, then row)
  int64_t seq;      // position in the direction's op order == summation order
};
struct Elem {
  int64_t buf, start, len;
  int32_t minStage;
};
} // namespace

static void buildFused(const BfCudaPlanDesc &desc, Direction &D, const std::vector<int32_t> &firstW,
                       const std::vector<int32_t> &lastW) {
  FusedTables &F = D.fused;
  F = FusedTables();
  for (const DirOp &d : D.ops)
    if (d.remote) return; // block-row shards exchange rows between stages: per-stage launches
  const int64_t nB = desc.numBufs;
  const int nQ = D.numQueues;
  const size_t nS = D.stages.size();
  // flat workspace: every written temp gets rows of its own (even count keeps real rows 16-byte aligned)
  std::vector<int64_t> flatOff((size_t)nB, -1);
  int64_t flatRows = 0;
  for (int64_t b = 0; b < nB; ++b) {
    if (b == D.inBufId || b == D.outBufId || lastW[(size_t)b] < 0) continue;
    flatOff[(size_t)b] = flatRows;
    flatRows += (desc.bufRows[b] + 1) & ~int64_t(1);
  }
  if (flatRows > (int64_t)kRowMask) return; // does not fit the 30-bit row addresses: keep per-stage launches
  F.wsRows = flatRows;
  (void)firstW;
  // counters.  Workspace buffer: one per buffer, bumped by every item that writes it.  Result (Y): one per
  // elementary interval that more than one stage writes (near-field rows + butterflies of different depth
  // accumulating into the same MatBlockDense row block); other result rows need none.
  std::vector<uint32_t> bufCounter((size_t)nB, kNoCounter), bufTotal((size_t)nB, 0);
  uint32_t nC = 0;
  for (int64_t b = 0; b < nB; ++b)
    if (flatOff[(size_t)b] >= 0) bufCounter[(size_t)b] = nC++;
  int64_t numElems = 0;
  for (const Stage &S : D.stages)
    for (int64_t e : S.itemElem) numElems = std::max(numElems, e + 1);
  std::vector<int32_t> elemFirstStage((size_t)numElems, -1);
  std::vector<uint8_t> elemMulti((size_t)numElems, 0);
  for (size_t s = 0; s < nS; ++s) {
    const Stage &S = D.stages[s];
    for (size_t i = 0; i < S.items.size(); ++i) {
      const int64_t b = S.itemBuf[i], e = S.itemElem[i];
      if (b == D.outBufId) {
        if (elemFirstStage[(size_t)e] < 0) elemFirstStage[(size_t)e] = (int32_t)s;
        else if (elemFirstStage[(size_t)e] != (int32_t)s) elemMulti[(size_t)e] = 1;
      } else {
        ++bufTotal[(size_t)b];
      }
    }
  }
  std::vector<uint32_t> elemCounter((size_t)numElems, kNoCounter);
  for (int64_t e = 0; e < numElems; ++e)
    if (elemMulti[(size_t)e]) elemCounter[(size_t)e] = nC++;
  F.numCounters = nC;
  // signals seen so far per counter, stage by stage: an accumulating item at stage s waits for the signals
  // of all earlier stages on its counter
  std::vector<uint32_t> before(nC, 0), during(nC, 0);
  std::vector<std::vector<ItemDep>> itemDeps(nS);
  for (size_t s = 0; s < nS; ++s) {
    const Stage &S = D.stages[s];
    itemDeps[s].resize(S.items.size());
    for (size_t i = 0; i < S.items.size(); ++i) {
      const int64_t b = S.itemBuf[i];
      const uint32_t c = b == D.outBufId ? elemCounter[(size_t)S.itemElem[i]] : bufCounter[(size_t)b];
      ItemDep dep{c, 0};
      if (c != kNoCounter) {
        if (S.items[i].flags & kItemAccumulate) dep.waitTarget = before[c];
        ++during[c];
      }
      itemDeps[s][i] = dep;
    }
    for (size_t i = 0; i < S.items.size(); ++i) {
      const uint32_t c = itemDeps[s][i].sig;
      if (c != kNoCounter && during[c]) { before[c] += during[c]; during[c] = 0; }
    }
  }
  auto flatAddr = [&](int64_t buf, int64_t row) -> uint32_t {
    if (buf == D.inBufId) return packAddr(kSpaceX, row);
    if (buf == D.outBufId) return packAddr(kSpaceY, row);
    return packAddr(kSpaceW, flatOff[(size_t)buf] + row);
  };
  // queue-major concatenation
  F.qItemBegin.assign((size_t)nQ + 1, 0);
  F.qTileBegin.assign((size_t)nQ + 1, 0);
  size_t totalItems = 0, totalTiles = 0;
  for (const Stage &S : D.stages) { totalItems += S.items.size(); totalTiles += S.tiles.size(); }
  F.items.reserve(totalItems + nS * (size_t)nQ / 2);
  F.itemDeps.reserve(totalItems + nS * (size_t)nQ / 2);
  F.tiles.reserve(totalTiles + nS * (size_t)nQ / 2);
  F.tileDeps.reserve(totalTiles + nS * (size_t)nQ / 2);
  for (int q = 0; q < nQ; ++q) {
    F.qItemBegin[(size_t)q] = (uint32_t)F.items.size();
    F.qTileBegin[(size_t)q] = (uint32_t)F.tiles.size();
    for (size_t s = 0; s < nS; ++s) {
      const Stage &S = D.stages[s];
      const uint32_t i0 = S.qItemBegin[(size_t)q], i1 = S.qItemBegin[(size_t)q + 1];
      if (i0 == i1) continue;
      { // the queue's pool span of this stage starts here
        const uint64_t p16 = S.qPool16[(size_t)q];
        DevItem ji{0u, (uint32_t)F.tiles.size(), 0, kItemJump, 1u};
        F.items.push_back(ji);
        F.itemDeps.push_back(ItemDep{kNoCounter, 0});
        F.tiles.push_back(DevTile{(uint32_t)(p16 & 0xFFFFFFFFull), kKcJump, (uint16_t)(p16 >> 32)});
        F.tileDeps.push_back(TileDep{kNoCounter, 0});
      }
      for (uint32_t i = i0; i < i1; ++i) {
        DevItem it = S.items[i];
        const uint32_t tb = it.tileBegin;
        it.outAddr = flatAddr(S.itemBuf[i], S.itemRow[i]);
        it.tileBegin = (uint32_t)F.tiles.size();
        F.items.push_back(it);
        F.itemDeps.push_back(itemDeps[s][i]);
        for (uint32_t t = tb; t < tb + it.numTiles; ++t) {
          DevTile tl = S.tiles[t];
          const int64_t b = S.tileBuf[t];
          tl.inAddr = flatAddr(b, S.tileRow[t]);
          F.tiles.push_back(tl);
          const uint32_t c = (b == D.inBufId) ? kNoCounter : bufCounter[(size_t)b];
          F.tileDeps.push_back(TileDep{c, c == kNoCounter ? 0u : bufTotal[(size_t)b]});
        }
      }
    }
  }
  F.qItemBegin[(size_t)nQ] = (uint32_t)F.items.size();
  F.qTileBegin[(size_t)nQ] = (uint32_t)F.tiles.size();
  if (F.tiles.size() > 0xFFFFFFF0ull || D.pool16 >> 48) { F = FusedTables(); return; }
  F.enabled = true;
}

// ---- step 0: stacks of small leaves -> one leaf ---------------------------------------------------------------
// The reference's streamed factorizations end in a block column of ~1000 leaves of 5-12 rows each over a shared
// column range (bfFacStreamer's Psi, src/fac.c:53-75; 2 120 leaves over 9 column ranges at 8 000 x 2 116).  Forward,
// each is a ragged item of its own; transposed, every one of them contributes a tile of 8 rows x ~7 columns to the
// same result rows: 352 000 tiles of 481 bytes, 0.08 of the HBM peak.  Leaves that are dense, not stored transposed,
// hold host values, share (inBuf, inOff, k, outBuf, flags) and follow one another in the output rows are therefore
// concatenated into one leaf (a copy of their values, owned by the Direction until the pool has been packed); the
// same for leaves side by side in the operand rows, and for runs of identity leaves.  Value-less descriptors
// (recordings, synthetic fills: values are defined per original leaf) and sharded ones are left alone.
static bool mergeStackedEnabled() { const char *e = std::getenv("BF_PLAN_MERGE"); return !e || std::atoi(e) != 0; }

static void mergeStackedLeaves(const BfCudaPlanDesc &desc, std::vector<BfCudaRawOp> &ops, std::vector<int64_t> &orig,
                               std::vector<std::vector<double>> &owned) {
  const size_t ed = desc.dtype == BF_CUDA_COMPLEX_F64 ? 2 : 1; // doubles per element
  constexpr int64_t kSmall = 64; // mean extent along the stacking axis below which a run is worth merging
  const int64_t n = (int64_t)ops.size();
  std::vector<uint8_t> dead((size_t)n, 0);
  auto mergeable = [&](const BfCudaRawOp &o) {
    if (o.m <= 0 || o.k <= 0 || o.flags != 0u) return false; // any flag (TRANS / CONJ / REMOTE / owner): leave alone
    if (o.kind == BF_CUDA_LEAF_IDENTITY) return o.m == o.k;
    return o.kind == BF_CUDA_LEAF_DENSE && o.val != nullptr && o.ld >= o.k;
  };
  // pass 0: vertical stacks (dense), pass 1: horizontal stacks (dense), pass 2: identity runs
  for (int pass = 0; pass < 3; ++pass) {
    std::vector<int64_t> idx;
    for (int64_t i = 0; i < n; ++i)
      if (!dead[(size_t)i] && mergeable(ops[(size_t)i]) && (ops[(size_t)i].kind == BF_CUDA_LEAF_IDENTITY) == (pass == 2)) idx.push_back(i);
    auto key = [&](int64_t i) {
      const BfCudaRawOp &o = ops[(size_t)i];
      if (pass == 0) return std::array<int64_t, 6>{o.inBuf, o.outBuf, o.inOff, o.k, o.outOff, i};
      if (pass == 1) return std::array<int64_t, 6>{o.inBuf, o.outBuf, o.outOff, o.m, o.inOff, i};
      return std::array<int64_t, 6>{o.inBuf, o.outBuf, o.outOff - o.inOff, 0, o.outOff, i};
    };
    std::sort(idx.begin(), idx.end(), [&](int64_t a, int64_t b) { return key(a) < key(b); });
    for (size_t a = 0; a < idx.size();) {
      size_t b = a + 1;
      auto follows = [&](int64_t p, int64_t q) { // q continues p
        const BfCudaRawOp &x = ops[(size_t)p], &y = ops[(size_t)q];
        if (x.inBuf != y.inBuf || x.outBuf != y.outBuf) return false;
        if (pass == 0) return x.inOff == y.inOff && x.k == y.k && x.outOff + x.m == y.outOff;
        if (pass == 1) return x.outOff == y.outOff && x.m == y.m && x.inOff + x.k == y.inOff;
        return x.outOff + x.m == y.outOff && x.inOff + x.k == y.inOff;
      };
      while (b < idx.size() && follows(idx[b - 1], idx[b])) ++b;
      const size_t cnt = b - a;
      int64_t ext = 0;
      for (size_t j = a; j < b; ++j) ext += pass == 1 ? ops[(size_t)idx[j]].k : ops[(size_t)idx[j]].m;
      if (cnt >= 2 && (pass == 2 || ext < kSmall * (int64_t)cnt)) {
        const int64_t first = *std::min_element(idx.begin() + (std::ptrdiff_t)a, idx.begin() + (std::ptrdiff_t)b);
        BfCudaRawOp M = ops[(size_t)idx[a]]; // lowest offset of the run
        if (pass == 2) {
          M.m = M.k = ext;
        } else {
          const int64_t rows = pass == 0 ? ext : M.m, cols = pass == 0 ? M.k : ext;
          std::vector<double> buf;
          try { buf.resize((size_t)rows * (size_t)cols * ed); } catch (const std::bad_alloc &) { a = b; continue; }
          int64_t r0 = 0, c0 = 0;
          for (size_t j = a; j < b; ++j) {
            const BfCudaRawOp &o = ops[(size_t)idx[j]];
            const double *src = static_cast<const double *>(o.val);
            for (int64_t r = 0; r < o.m; ++r)
              std::memcpy(buf.data() + ((size_t)(r0 + r) * (size_t)cols + (size_t)c0) * ed, src + (size_t)r * (size_t)o.ld * ed, (size_t)o.k * ed * sizeof(double));
            if (pass == 0) r0 += o.m; else c0 += o.k;
          }
          M.m = rows; M.k = cols; M.ld = cols;
          owned.push_back(std::move(buf));
          M.val = owned.back().data();
        }
        for (size_t j = a; j < b; ++j) dead[(size_t)idx[j]] = 1;
        dead[(size_t)first] = 0;
        ops[(size_t)first] = M; // (orig[first] stays: the run is known by its first leaf)
      }
      a = b;
    }
  }
  size_t w = 0;
  for (int64_t i = 0; i < n; ++i)
    if (!dead[(size_t)i]) { ops[w] = ops[(size_t)i]; orig[w] = orig[(size_t)i]; ++w; }
  ops.resize(w);
  orig.resize(w);
}

// An identity leaf that is the ONLY writer of a whole temporary B (B = A[off : off + m]) is a copy: whoever reads B
// reads A instead.  One that is the only READER of a whole temporary A is a copy too: whoever writes A writes its
// rows of the identity's output instead (other writers of those rows simply accumulate, as they did).  The streamed
// factorizations of the reference carry such a factor at their row end (after step 0 it is ONE identity leaf over
// all rows): a stage and a launch in each direction for nothing but a copy of the result / operand.
static bool elideIdentityEnabled() { const char *e = std::getenv("BF_PLAN_ELIDE"); return !e || std::atoi(e) != 0; }

static void elideIdentityCopies(const BfCudaPlanDesc &desc, std::vector<BfCudaRawOp> &ops, std::vector<int64_t> &orig) {
  const int64_t nB = desc.numBufs;
  for (bool again = true; again;) {
    again = false;
    std::vector<int64_t> writers((size_t)nB, 0), readers((size_t)nB, 0);
    for (const BfCudaRawOp &o : ops)
      if (o.m > 0 && o.k > 0) { ++writers[(size_t)o.outBuf]; ++readers[(size_t)o.inBuf]; }
    for (size_t i = 0; i < ops.size(); ++i) {
      const BfCudaRawOp I = ops[i];
      if (I.kind != BF_CUDA_LEAF_IDENTITY || I.flags != 0u || I.m != I.k || I.m <= 0 || I.inBuf == I.outBuf) continue;
      const int64_t A = I.inBuf, B = I.outBuf;
      const bool tempA = A != BF_CUDA_BUF_X && A != BF_CUDA_BUF_Y, tempB = B != BF_CUDA_BUF_X && B != BF_CUDA_BUF_Y;
      bool done = false;
      if (tempB && writers[(size_t)B] == 1 && I.outOff == 0 && I.m == desc.bufRows[B]) {
        bool ok = true; // (a reader of B must not end up reading the buffer it writes)
        for (const BfCudaRawOp &o : ops) if (o.inBuf == B && o.outBuf == A) ok = false;
        if (ok) {
          for (BfCudaRawOp &o : ops) if (o.inBuf == B) { o.inBuf = A; o.inOff += I.inOff; }
          done = true;
        }
      }
      if (!done && tempA && readers[(size_t)A] == 1 && I.inOff == 0 && I.m == desc.bufRows[A]) {
        bool ok = true;
        for (const BfCudaRawOp &o : ops) if (o.outBuf == A && o.inBuf == B) ok = false;
        if (ok) {
          for (BfCudaRawOp &o : ops) if (o.outBuf == A) { o.outBuf = B; o.outOff += I.outOff; }
          done = true;
        }
      }
      if (done) {
        ops.erase(ops.begin() + (std::ptrdiff_t)i);
        orig.erase(orig.begin() + (std::ptrdiff_t)i);
        again = true;
        break;
      }
    }
  }
}

BfCudaStatus compileDirection(const BfCudaPlanDesc &desc, int32_t op, int numQueues, Direction &D,
                              std::string &err) {
  const bool cplx = desc.dtype == BF_CUDA_COMPLEX_F64;
  const bool rev = op != BF_CUDA_OP_N;
  if (op == BF_CUDA_OP_H && !cplx) { err = "OP_H on a real plan"; return BF_CUDA_ERROR_INVALID_ARGUMENT; }
  if (desc.numBufs < 2 || !desc.bufRows || (desc.numOps > 0 && !desc.ops)) {
    err = "malformed descriptor"; return BF_CUDA_ERROR_INVALID_ARGUMENT;
  }
  if (desc.bufRows[BF_CUDA_BUF_X] != desc.numCols || desc.bufRows[BF_CUDA_BUF_Y] != desc.numRows) {
    err = "buffer 0 / 1 must have numCols / numRows rows"; return BF_CUDA_ERROR_INVALID_ARGUMENT;
  }
  const bool sharded = desc.shardWorld > 1;
  if (sharded && (desc.shardRank < 0 || desc.shardRank >= desc.shardWorld || desc.shardWorld > 16)) {
    err = "shardRank / shardWorld out of range"; return BF_CUDA_ERROR_INVALID_ARGUMENT;
  }
  if (sharded && rev) { err = "block-row shards are forward only"; return BF_CUDA_ERROR_NOT_IMPLEMENTED; }
  D = Direction();
  D.op = op;
  D.dtype = desc.dtype;
  D.numQueues = numQueues;
  D.shardWorld = sharded ? desc.shardWorld : 0;
  D.shardRank = sharded ? desc.shardRank : 0;
  D.inBufId = rev ? BF_CUDA_BUF_Y : BF_CUDA_BUF_X;
  D.outBufId = rev ? BF_CUDA_BUF_X : BF_CUDA_BUF_Y;
  D.rowsIn = desc.bufRows[D.inBufId];
  D.rowsOut = desc.bufRows[D.outBufId];
  const int64_t nB = desc.numBufs;
  // ---- 0. stacks of small leaves become one leaf (see mergeStackedLeaves) ----------------------
  std::vector<BfCudaRawOp> rawOps(desc.ops, desc.ops + desc.numOps);
  std::vector<int64_t> rawOrig((size_t)desc.numOps);
  std::iota(rawOrig.begin(), rawOrig.end(), int64_t(0));
  if (!sharded && !(desc.buildTranspose & BF_CUDA_BUILD_KEEP_LEAVES) && mergeStackedEnabled()) {
    bool sane = true; // (ranges are validated below, on whatever list we end up with; merge only what is sane)
    for (const BfCudaRawOp &o : rawOps)
      if ((o.flags & ~(BF_CUDA_LEAF_TRANS | BF_CUDA_LEAF_CONJ)) != 0u || // REMOTE / owner bits: somebody shards by leaf and buffer
          o.inBuf < 0 || o.inBuf >= nB || o.outBuf < 0 || o.outBuf >= nB || o.m < 0 || o.k < 0 || o.inOff < 0 || o.outOff < 0 ||
          o.inOff + o.k > desc.bufRows[o.inBuf] || o.outOff + o.m > desc.bufRows[o.outBuf]) { sane = false; break; }
    if (sane) mergeStackedLeaves(desc, rawOps, rawOrig, D.ownedValues);
    if (sane && elideIdentityEnabled()) elideIdentityCopies(desc, rawOps, rawOrig);
  }
  const int64_t nOps = (int64_t)rawOps.size();

  // ---- 1. leaves as they act in this direction, in processing order -------------------------
  D.ops.reserve((size_t)nOps);
  for (int64_t n = 0; n < nOps; ++n) {
    const int64_t rn = rev ? nOps - 1 - n : n;
    const int64_t ri = rawOrig[(size_t)rn]; // index in the caller's list (of the first leaf of a merged run)
    const BfCudaRawOp &o = rawOps[(size_t)rn];
    if (o.inBuf < 0 || o.inBuf >= nB || o.outBuf < 0 || o.outBuf >= nB || o.m < 0 || o.k < 0 ||
        o.inOff < 0 || o.outOff < 0 || o.inOff + o.k > desc.bufRows[o.inBuf] ||
        o.outOff + o.m > desc.bufRows[o.outBuf] || o.inBuf == o.outBuf) {
      err = "op " + std::to_string(ri) + ": range outside its buffers"; return BF_CUDA_ERROR_INVALID_ARGUMENT;
    }
    if (o.kind != BF_CUDA_LEAF_DENSE && o.kind != BF_CUDA_LEAF_IDENTITY && o.kind != BF_CUDA_LEAF_DIAG) {
      err = "op " + std::to_string(ri) + ": unknown leaf kind"; return BF_CUDA_ERROR_NOT_IMPLEMENTED;
    }
    if (o.m == 0 || o.k == 0) continue;
    DirOp d{};
    d.kind = o.kind;
    d.remote = (o.flags & BF_CUDA_LEAF_REMOTE) ? 1 : 0;
    d.owner = (uint8_t)(sharded ? BF_CUDA_LEAF_OWNER(o.flags) : 0);
    if (sharded && ((int)d.owner >= desc.shardWorld || (d.remote != 0) != ((int)d.owner != desc.shardRank))) {
      err = "op " + std::to_string(ri) + ": owner bits and REMOTE flag disagree"; return BF_CUDA_ERROR_INVALID_ARGUMENT;
    }
    d.rawIndex = ri;
    d.val = o.val;
    d.ld = o.ld;
    const bool st = o.kind == BF_CUDA_LEAF_DENSE && leafStoredTransposed(o, cplx);
    // acting(a,b) = st ? conj(stored(b,a)) : stored(a,b);  N: D = acting, T: D = acting^T, H: acting^H
    d.swapRC = (uint8_t)(st ^ rev);
    d.conj = (uint8_t)(cplx && (st ^ (op == BF_CUDA_OP_H)));
    d.storedCols = st ? o.m : o.k;
    if (o.kind == BF_CUDA_LEAF_DENSE && o.ld < d.storedCols) {
      err = "op " + std::to_string(ri) + ": ld smaller than the stored column count"; return BF_CUDA_ERROR_INVALID_ARGUMENT;
    }
    d.inBuf = rev ? o.outBuf : o.inBuf;  d.inOff = rev ? o.outOff : o.inOff;
    d.outBuf = rev ? o.inBuf : o.outBuf; d.outOff = rev ? o.inOff : o.outOff;
    d.m = rev ? o.k : o.m;               d.k = rev ? o.m : o.k;
    if (o.kind != BF_CUDA_LEAF_DENSE) { // identity / diagonal act on min(m,k) rows; the rest is zero
      d.m = d.k = std::min(o.m, o.k);
      d.swapRC = 0; d.conj = 0; d.storedCols = 1;
    }
    D.ops.push_back(d);
  }

  // ---- 2. stages: an op runs once every writer of its operand buffer has run ------------------
  std::vector<int32_t> level((size_t)nB, 0), firstW((size_t)nB, INT_MAX), lastW((size_t)nB, -1);
  std::vector<uint8_t> wasRead((size_t)nB, 0);
  int32_t numStages = 1;
  for (DirOp &d : D.ops) {
    if (d.outBuf == D.inBufId) { err = "an op writes the operand buffer"; return BF_CUDA_ERROR_INVALID_ARGUMENT; }
    if (d.inBuf == D.outBufId) { err = "an op reads the result buffer"; return BF_CUDA_ERROR_INVALID_ARGUMENT; }
    if (wasRead[d.outBuf]) { err = "buffer " + std::to_string(d.outBuf) + " written after being read (ops not in dependency order)"; return BF_CUDA_ERROR_INVALID_ARGUMENT; }
    wasRead[d.inBuf] = 1;
    d.stage = level[d.inBuf];
    level[d.outBuf] = std::max(level[d.outBuf], d.stage + 1);
    firstW[d.outBuf] = std::min(firstW[d.outBuf], d.stage);
    lastW[d.outBuf] = std::max(lastW[d.outBuf], d.stage);
    numStages = std::max(numStages, d.stage + 1);
    if (d.kind == BF_CUDA_LEAF_DENSE && !d.remote) D.entries += d.m * d.k;
  }
  for (const DirOp &d : D.ops)
    if (d.inBuf != D.inBufId && lastW[d.inBuf] < 0) {
      err = "an op reads buffer " + std::to_string(d.inBuf) + " which nothing writes"; return BF_CUDA_ERROR_INVALID_ARGUMENT;
    }
  // (Tried and measured slower, so NOT done: dealing the leaves that read X and write Y — which depend on nothing —
  // out over the later stages so that every stage carries the same bytes.  N = 16 384: 0.142 -> 0.152 ms, N = 65 536:
  // 0.578 -> 0.616 ms; the extra, smaller items on Y cost more than the short stages do.  profiles/r2i_plan_knobs.txt)
  D.stages.assign((size_t)numStages, Stage());

  // ---- 3. workspace: ping-pong arenas keyed by the parity of the stage that READS a buffer ----
  // A temp written only at stage s-1 and read at stage s is live during s-1 and s; temps read at
  // s+1 are born at s.  Two arenas (even / odd read stage) therefore never alias live data.
  // Buffers written over several stages (a MatProduct factor that itself contains products) get
  // dedicated rows.
  D.wsOffset.assign((size_t)nB, -1);
  {
    std::vector<int64_t> perStage((size_t)numStages + 2, 0);
    int64_t dedicated = 0;
    std::vector<int64_t> off((size_t)nB, 0);
    std::vector<int8_t> cls((size_t)nB, -1); // 0/1 arena parity, 2 dedicated, -1 unused
    for (int64_t b = 0; b < nB; ++b) {
      if (b == D.inBufId || b == D.outBufId) continue;
      if (lastW[b] < 0) continue; // never written: never read either unless it is all zeros
      const int64_t rows = (desc.bufRows[b] + 1) & ~int64_t(1); // even, keeps 16-byte alignment for real
      if (firstW[b] == lastW[b]) {
        const int32_t rs = level[b];
        cls[b] = (int8_t)(rs & 1);
        off[b] = perStage[rs];
        perStage[rs] += rows;
      } else {
        cls[b] = 2;
        off[b] = dedicated;
        dedicated += rows;
      }
    }
    int64_t arena[2] = {0, 0};
    if (sharded) {
      // peers store into this workspace while it is being read: every buffer keeps rows of its own
      int64_t flat = 0;
      for (int64_t b = 0; b < nB; ++b)
        if (cls[b] >= 0) { D.wsOffset[b] = flat; flat += (desc.bufRows[b] + 1) & ~int64_t(1); }
      D.wsRows = flat;
    } else {
      for (int32_t s = 0; s < numStages + 2; ++s) arena[s & 1] = std::max(arena[s & 1], perStage[s]);
      for (int64_t b = 0; b < nB; ++b) {
        if (cls[b] == 0) D.wsOffset[b] = off[b];
        else if (cls[b] == 1) D.wsOffset[b] = arena[0] + off[b];
        else if (cls[b] == 2) D.wsOffset[b] = arena[0] + arena[1] + off[b];
      }
      D.wsRows = arena[0] + arena[1] + dedicated;
    }
    D.wsArena[0] = arena[0]; D.wsArena[1] = arena[1];
    if (D.wsRows > (int64_t)kRowMask || D.rowsIn > (int64_t)kRowMask || D.rowsOut > (int64_t)kRowMask) {
      err = "operator too large for 30-bit row addresses"; return BF_CUDA_ERROR_NOT_IMPLEMENTED;
    }
  }
  auto addrOf = [&](int64_t buf, int64_t row) -> uint32_t {
    if (buf == D.inBufId) return packAddr(kSpaceX, row);
    if (buf == D.outBufId) return packAddr(kSpaceY, row);
    return packAddr(kSpaceW, D.wsOffset[buf] + row);
  };

  // ---- 4. elementary output intervals and (leaf, interval) incidences -------------------------
  // Every output buffer is cut at all outOff / outOff + m of the leaves that write it; a leaf
  // covers a run of whole elementary intervals, and rows of a row-major leaf stay contiguous
  // under row splitting.  An item is an elementary interval (cut again into <= R rows) at one
  // stage; it stores (first stage to touch those rows) or accumulates (later stages).
  const int64_t nD = (int64_t)D.ops.size();
  std::vector<int64_t> byOut((size_t)nD);
  std::iota(byOut.begin(), byOut.end(), int64_t(0));
  std::sort(byOut.begin(), byOut.end(), [&](int64_t a, int64_t b) {
    const DirOp &x = D.ops[a], &y = D.ops[b];
    if (x.outBuf != y.outBuf) return x.outBuf < y.outBuf;
    if (x.outOff != y.outOff) return x.outOff < y.outOff;
    return a < b;
  });
  std::vector<Elem> elems;
  std::vector<uint8_t> remoteCovered; // per elem
  std::vector<Frag> frags;
  frags.reserve((size_t)nD + 16);
  std::vector<int64_t> bps;
  auto pushGap = [&](int64_t buf, int64_t a, int64_t b) { // rows [a,b) of buf that no leaf writes
    if (b <= a) return;
    const int32_t st = firstW[buf] == INT_MAX ? 0 : firstW[buf];
    elems.push_back(Elem{buf, a, b - a, st});
    frags.push_back(Frag{st, (int64_t)elems.size() - 1, -1});
  };
  {
    int64_t g = 0;
    std::vector<uint8_t> seen((size_t)nB, 0);
    while (g < nD) {
      const int64_t buf = D.ops[byOut[g]].outBuf;
      int64_t h = g;
      bps.clear();
      while (h < nD && D.ops[byOut[h]].outBuf == buf) {
        const DirOp &d = D.ops[byOut[h]];
        bps.push_back(d.outOff);
        bps.push_back(d.outOff + d.m);
        ++h;
      }
      std::sort(bps.begin(), bps.end());
      bps.erase(std::unique(bps.begin(), bps.end()), bps.end());
      seen[buf] = 1;
      pushGap(buf, 0, bps.front());
      const int64_t e0 = (int64_t)elems.size();
      for (size_t i = 0; i + 1 < bps.size(); ++i) elems.push_back(Elem{buf, bps[i], bps[i + 1] - bps[i], INT_MAX});
      for (int64_t i = g; i < h; ++i) {
        const DirOp &d = D.ops[byOut[i]];
        const int64_t lo = std::lower_bound(bps.begin(), bps.end(), d.outOff) - bps.begin();
        const int64_t hi = std::lower_bound(bps.begin(), bps.end(), d.outOff + d.m) - bps.begin();
        for (int64_t e = lo; e < hi; ++e) {
          if (d.remote) { // another rank produces these rows: no items here, and not a hole either
            if ((int64_t)remoteCovered.size() <= e0 + e) remoteCovered.resize((size_t)(e0 + e) + 1, 0);
            remoteCovered[(size_t)(e0 + e)] = 1;
            continue;
          }
          frags.push_back(Frag{d.stage, e0 + e, byOut[i]});
          elems[(size_t)(e0 + e)].minStage = std::min(elems[(size_t)(e0 + e)].minStage, d.stage);
        }
      }
      const int64_t e1 = (int64_t)elems.size();
      for (int64_t e = e0; e < e1; ++e)
        if (elems[(size_t)e].minStage == INT_MAX && !((int64_t)remoteCovered.size() > e && remoteCovered[(size_t)e])) { // hole between two leaves
          elems[(size_t)e].minStage = firstW[buf];
          frags.push_back(Frag{firstW[buf], e, -1});
        }
      pushGap(buf, bps.back(), desc.bufRows[buf]);
      g = h;
    }
    // nothing writes the result at all: it is all zeros
    for (int64_t b = 0; b < nB; ++b)
      if (!seen[b] && b == D.outBufId) pushGap(b, 0, desc.bufRows[b]);
  }
  std::sort(frags.begin(), frags.end(), [](const Frag &a, const Frag &b) {
    if (a.stage != b.stage) return a.stage < b.stage;
    if (a.elem != b.elem) return a.elem < b.elem;
    return a.seq < b.seq;
  });

  // ---- 4b. result rows too long for one queue ------------------------------------------------------
  // An item is at least 8 rows of an elementary interval with ALL the leaves that write those rows at that stage,
  // and it is read by one warp.  On a small operator the near-field leaves of one block row (X -> Y, stage 0) can
  // exceed a queue's share of the stage even at 8 rows (N = 16 384: 283 KB against 174 KB, 48 such items): every
  // other queue then waits for them (SM sub-partitions active 63 % of stage 0, ncu r2y).  Leaves that read the
  // operand and write the result depend on nothing and may run at ANY stage (later stages accumulate), so the
  // excess is handed, leaf by leaf, to the later stage where the same rows are lightest relative to that stage's
  // share.  (Dealing ALL such leaves out evenly was measured slower — see step 2; this moves a few percent of
  // stage 0.)  Only the frag moves: the summation order of those rows changes, nothing else.
  if (!sharded && numStages > 1 && deferHeavyEnabled()) {
    const int64_t es = desc.dtype == BF_CUDA_COMPLEX_F64 ? 16 : 8;
    std::vector<int64_t> sBytes((size_t)numStages, 0);
    std::map<std::pair<int32_t, int64_t>, int64_t> rowBytesOf; // (stage, elem) -> bytes per row
    for (const Frag &fr : frags)
      if (fr.seq >= 0 && D.ops[(size_t)fr.seq].kind == BF_CUDA_LEAF_DENSE) {
        sBytes[(size_t)fr.stage] += elems[(size_t)fr.elem].len * D.ops[(size_t)fr.seq].k * es;
        rowBytesOf[{fr.stage, fr.elem}] += D.ops[(size_t)fr.seq].k * es;
      }
    std::vector<double> share((size_t)numStages);
    for (int32_t t = 0; t < numStages; ++t) share[(size_t)t] = std::max<double>(1.0, (double)sBytes[(size_t)t] / std::max(numQueues, 1));
    bool moved = false;
    for (size_t f = 0; f < frags.size();) {
      size_t f1 = f;
      while (f1 < frags.size() && frags[f1].stage == frags[f].stage && frags[f1].elem == frags[f].elem) ++f1;
      const Elem &el = elems[(size_t)frags[f].elem];
      const int32_t s = frags[f].stage;
      const size_t g0 = f, g1 = f1;
      f = f1;
      if (el.buf != D.outBufId || frags[g0].seq < 0 || s + 1 >= numStages) continue;
      const double rowsMin = (double)std::min<int64_t>(el.len, 8);
      int64_t kept = rowBytesOf[{s, frags[g0].elem}];
      if (rowsMin * (double)kept <= 1.1 * share[(size_t)s]) continue;
      size_t left = g1 - g0;
      for (size_t q = g1; q-- > g0 && left > 1;) {
        const DirOp &d = D.ops[(size_t)frags[q].seq];
        if (d.kind != BF_CUDA_LEAF_DENSE || d.inBuf != D.inBufId) continue;
        if (rowsMin * (double)kept <= share[(size_t)s]) break;
        const int64_t rb = d.k * es;
        int32_t best = -1;
        double bestRatio = rowsMin * (double)kept / share[(size_t)s]; // moving must leave the rows lighter there than here
        for (int32_t t = s + 1; t < numStages; ++t) {
          auto itT = rowBytesOf.find({t, frags[q].elem});
          const double have = itT == rowBytesOf.end() ? 0.0 : (double)itT->second;
          const double ratio = rowsMin * (have + (double)rb) / share[(size_t)t];
          if (ratio < bestRatio) { bestRatio = ratio; best = t; }
        }
        if (best < 0) continue;
        rowBytesOf[{best, frags[q].elem}] += rb;
        kept -= rb;
        frags[q].stage = best;
        --left;
        moved = true;
      }
    }
    if (moved)
      std::sort(frags.begin(), frags.end(), [](const Frag &a, const Frag &b) {
        if (a.stage != b.stage) return a.stage < b.stage;
        if (a.elem != b.elem) return a.elem < b.elem;
        return a.seq < b.seq;
      });
  }

  // ---- 5. items and tiles per stage -------------------------------------------------------------
  // An item is read by ONE warp (1-RHS kernel), so an item much longer than a queue's share of its
  // stage is a tail nobody can help with.  Operators that fill the GPU many times over never see
  // this; on small ones (N ~ 16k: 170 KB per queue, but 32 rows of a first-level block are 700 KB)
  // it tripled the stage time.  Items whose rows are long are therefore cut into fewer rows (down
  // to 8, one DMMA fragment) so that an item stays near the per-queue share.
  const int Rmax = maxItemRows(desc.dtype);
  const int64_t esBytes = desc.dtype == BF_CUDA_COMPLEX_F64 ? 16 : 8;
  std::vector<int64_t> stageBytes(D.stages.size(), 0);
  for (const Frag &fr : frags)
    if (fr.seq >= 0 && D.ops[(size_t)fr.seq].kind == BF_CUDA_LEAF_DENSE)
      stageBytes[(size_t)fr.stage] += elems[(size_t)fr.elem].len * D.ops[(size_t)fr.seq].k * esBytes;
  for (size_t f = 0; f < frags.size();) {
    size_t f1 = f;
    while (f1 < frags.size() && frags[f1].stage == frags[f].stage && frags[f1].elem == frags[f].elem) ++f1;
    const Elem &el = elems[(size_t)frags[f].elem];
    Stage &S = D.stages[(size_t)frags[f].stage];
    const bool zeroOnly = frags[f].seq < 0;
    int R = Rmax;
    if (!zeroOnly && el.len > 8) {
      int64_t rowBytes = 0;
      for (size_t q = f; q < f1; ++q)
        if (D.ops[(size_t)frags[q].seq].kind == BF_CUDA_LEAF_DENSE) rowBytes += D.ops[(size_t)frags[q].seq].k * esBytes;
      // (not finer than that: cutting down to a quarter of the share evens the queues out on paper, but 8-row items
      // move 4 KiB tiles and the apply got 10 % SLOWER at N = 65 536 — measured, profiles/r2i_plan_knobs.txt)
      const int64_t share = stageBytes[(size_t)frags[f].stage] / std::max(numQueues, 1);
      while (R > 8 && (int64_t)R * rowBytes > share) R -= 8;
    }
    const int64_t nChunks = (el.len + R - 1) / R;
    for (int64_t c = 0; c < nChunks; ++c) {
      // cut at multiples of 8 rows (whole m8 DMMA fragments); only the last chunk is ragged
      const int64_t r0 = std::min(el.len, (el.len * c / nChunks + 7) & ~int64_t(7));
      const int64_t r1 = c + 1 == nChunks ? el.len : std::min(el.len, (el.len * (c + 1) / nChunks + 7) & ~int64_t(7));
      const int rows = (int)(r1 - r0);
      DevItem it{};
      it.outAddr = addrOf(el.buf, el.start + r0);
      it.tileBegin = (uint32_t)S.tiles.size();
      it.rows = (uint16_t)rows;
      it.flags = (uint16_t)((!zeroOnly && frags[f].stage > el.minStage) ? kItemAccumulate : 0);
      if (!zeroOnly) {
        const int KC = kcForRows(desc.dtype, rows);
        for (size_t q = f; q < f1; ++q) {
          const DirOp &d = D.ops[(size_t)frags[q].seq];
          const int64_t rowInOp = el.start + r0 - d.outOff;
          if (d.kind != BF_CUDA_LEAF_DENSE) {
            DevTile t{addrOf(d.inBuf, d.inOff + rowInOp), d.kind == BF_CUDA_LEAF_DIAG ? kKcDiag : kKcIdentity, (uint16_t)rows};
            S.tiles.push_back(t);
            S.pieces.push_back(PieceRef{frags[q].seq, (int32_t)rowInOp, (int32_t)rowInOp});
            S.tileBuf.push_back(d.inBuf); S.tileRow.push_back(d.inOff + rowInOp);
            continue;
          }
          const int64_t nK = (d.k + KC - 1) / KC;
          for (int64_t kk = 0; kk < nK; ++kk) {
            // pieces of whole k4 DMMA steps; only the last one is ragged
            const int64_t c0 = std::min(d.k, (d.k * kk / nK + 3) & ~int64_t(3));
            const int64_t c1 = kk + 1 == nK ? d.k : std::min(d.k, (d.k * (kk + 1) / nK + 3) & ~int64_t(3));
            DevTile t{addrOf(d.inBuf, d.inOff + c0), (uint16_t)(c1 - c0), (uint16_t)rows};
            S.tiles.push_back(t);
            S.pieces.push_back(PieceRef{frags[q].seq, (int32_t)rowInOp, (int32_t)c0});
            S.tileBuf.push_back(d.inBuf); S.tileRow.push_back(d.inOff + c0);
            D.maxTileK = std::max<int64_t>(D.maxTileK, c1 - c0);
          }
          S.entries += (int64_t)rows * d.k;
        }
      }
      it.numTiles = (uint32_t)S.tiles.size() - it.tileBegin;
      S.items.push_back(it);
      S.itemBuf.push_back(el.buf); S.itemElem.push_back(frags[f].elem); S.itemRow.push_back(el.start + r0);
      D.maxItemRows = std::max<int64_t>(D.maxItemRows, rows);
    }
    f = f1;
  }

  // ---- 6. static work queues + pool layout -------------------------------------------------------
  // Items of a stage are dealt, in order, to numQueues queues of nearly equal cost (bytes streamed
  // plus a fixed charge per tile and per item).  The pool is laid out in that same order, so queue
  // q streams the contiguous span [qPool16[q], qPool16[q+1]) of HBM.
  uint64_t pool16 = 0;
  for (Stage &S : D.stages) {
    const bool lpt = sortItemsEnabled();
    std::vector<uint32_t> lptEnd;
    if (lpt) lptEnd = assignItemsLpt(desc.dtype, S, numQueues);
    const size_t nI = S.items.size(), nT = S.tiles.size();
    S.tilePool16.resize(nT);
    S.poolBegin16 = pool16;
    std::vector<uint64_t> cost(nI + 1, 0);
    for (size_t i = 0; i < nI; ++i) {
      const DevItem &it = S.items[i];
      uint64_t c = 256 + 32ull * it.rows;
      for (uint32_t t = it.tileBegin; t < it.tileBegin + it.numTiles; ++t) {
        S.tilePool16[t] = pool16;
        const uint64_t b16 = tileBytes16(desc.dtype, S.tiles[t].rows, S.tiles[t].kc);
        pool16 += b16;
        c += 16 * b16 + 384;
      }
      cost[i + 1] = cost[i] + c;
    }
    S.poolEnd16 = pool16;
    S.qItemBegin.assign((size_t)numQueues + 1, (uint32_t)nI);
    S.qTileBegin.assign((size_t)numQueues + 1, (uint32_t)nT);
    S.qPool16.assign((size_t)numQueues + 1, pool16);
    // Contiguous partition with the smallest possible bottleneck: the least capacity C for which "fill a queue
    // until the next item would overflow C" needs at most numQueues queues (binary search), then a pass that stops
    // a queue early once it holds its share of what is left (even queues instead of full ones followed by empty
    // ones), falling back to the plain fill if that runs out of queues.  A stage lasts as long as its slowest
    // queue; the former rule (deal items while their midpoint lies before the queue's share boundary) left some
    // queues with two share-sized items: slowest / mean queue bytes 1.9 on an 8-way row shard of N = 524 288, now
    // 1.2 - 1.35; N = 16 384 apply 0.153 -> 0.142 ms, 8-GPU shard 1.86 -> 1.77 ms.
    const uint64_t total = cost[nI];
    std::vector<uint32_t> qEnd((size_t)numQueues, (uint32_t)nI); // items [qEnd[q-1], qEnd[q]) belong to queue q
    if (lpt) {
      qEnd = lptEnd;
    } else if (nI > 0) {
      uint64_t lo = 0;
      for (size_t i = 0; i < nI; ++i) lo = std::max(lo, cost[i + 1] - cost[i]);
      uint64_t hi = total;
      lo = std::max(lo, (total + (uint64_t)numQueues - 1) / (uint64_t)numQueues);
      auto queuesNeeded = [&](uint64_t C) {
        int used = 1;
        uint64_t acc = 0;
        for (size_t i = 0; i < nI; ++i) {
          const uint64_t c = cost[i + 1] - cost[i];
          if (acc + c > C && acc > 0) { ++used; acc = 0; }
          acc += c;
        }
        return used;
      };
      while (lo < hi) {
        const uint64_t mid = lo + (hi - lo) / 2;
        if (queuesNeeded(mid) <= numQueues) hi = mid; else lo = mid + 1;
      }
      const uint64_t C = lo;
      auto fill = [&](bool even) {
        size_t i = 0;
        for (int q = 0; q < numQueues; ++q) {
          const uint64_t left = total - cost[i];
          const uint64_t fair = (left + (uint64_t)(numQueues - q) - 1) / (uint64_t)(numQueues - q);
          uint64_t acc = 0;
          while (i < nI) {
            const uint64_t c = cost[i + 1] - cost[i];
            if (acc > 0 && acc + c > C) break;
            if (even && acc > 0 && acc + c / 2 > fair) break;
            acc += c;
            ++i;
          }
          qEnd[(size_t)q] = (uint32_t)i;
        }
        return i == nI;
      };
      if (!fill(true)) fill(false);
      qEnd[(size_t)numQueues - 1] = (uint32_t)nI;
    }
    for (int q = 0; q < numQueues; ++q) {
      const size_t i = q == 0 ? 0 : qEnd[(size_t)q - 1];
      S.qItemBegin[(size_t)q] = (uint32_t)i;
      S.qTileBegin[(size_t)q] = i < nI ? S.items[i].tileBegin : (uint32_t)nT;
      S.qPool16[(size_t)q] = (i < nI && S.items[i].tileBegin < nT) ? S.tilePool16[S.items[i].tileBegin] : pool16;
    }
    // fix pool offsets of queues that begin at an item without tiles: the next tile in order
    for (int q = numQueues - 1; q >= 0; --q) {
      const uint32_t tb = S.qTileBegin[(size_t)q];
      S.qPool16[(size_t)q] = tb < nT ? S.tilePool16[tb] : pool16;
    }
    D.numItems += (int64_t)nI;
    D.numTiles += (int64_t)nT;
  }
  D.pool16 = pool16;

  // ---- 7. fused (single-launch) tables: queue-major, flat workspace, readiness counters ------------
  buildFused(desc, D, firstW, lastW);

  // ---- 8. block-row shards: which of my rows which peer reads ---------------------------------------
  if (sharded) {
    const uint32_t me = 1u << desc.shardRank;
    std::vector<uint16_t> readers((size_t)D.wsRows + 1, 0); // per workspace row: ranks whose leaves read it
    for (const DirOp &d : D.ops)
      if (d.inBuf != D.inBufId) {
        uint16_t *r = readers.data() + D.wsOffset[(size_t)d.inBuf] + d.inOff;
        for (int64_t i = 0; i < d.k; ++i) r[i] |= (uint16_t)(1u << d.owner);
      }
    for (size_t s = 0; s < D.stages.size(); ++s) {
      Stage &S = D.stages[s];
      S.itemPeers.assign(S.items.size(), 0u);
      for (size_t i = 0; i < S.items.size(); ++i) {
        const DevItem &it = S.items[i];
        if ((it.outAddr >> 30) != kSpaceW) continue;
        uint32_t m = 0;
        const uint16_t *r = readers.data() + (it.outAddr & kRowMask);
        for (int q = 0; q < it.rows; ++q) m |= r[q];
        m &= ~me;
        S.itemPeers[i] = m;
        S.sendMask |= m;
        for (uint32_t b = m; b; b &= b - 1) S.sentRows += it.rows;
      }
    }
    for (const DirOp &d : D.ops)
      if (d.remote && d.outBuf != D.outBufId) {
        const uint16_t *r = readers.data() + D.wsOffset[(size_t)d.outBuf] + d.outOff;
        bool mine = false;
        for (int64_t i = 0; i < d.m && !mine; ++i) mine = (r[i] & me) != 0;
        if (mine) D.stages[(size_t)d.stage].recvMask |= 1u << d.owner;
      }
  }
  return BF_CUDA_OK;
}

// ---- value packing ------------------------------------------------------------------------------
template <typename T> static inline T conjIf(T v, bool) { return v; }
template <> inline std::complex<double> conjIf(std::complex<double> v, bool c) { return c ? std::conj(v) : v; }

// tiles [t0, t1) of stage S; base[0] corresponds to pool offset base16
template <typename T>
static void packStage(const Direction &dir, const Stage &S, T *base, size_t t0, size_t t1, uint64_t base16) {
  for (size_t t = t0; t < t1; ++t) {
    const DevTile &tl = S.tiles[t];
    if (tl.kc == kKcIdentity) continue;
    const PieceRef &pr = S.pieces[t];
    const DirOp &d = dir.ops[(size_t)pr.op];
    if (!d.val) continue;
    char *dstBytes = reinterpret_cast<char *>(base) + 16 * (S.tilePool16[t] - base16);
    if (tl.kc == kKcDiag) { // real diagonal values, also on a complex operator (src/mat_diag_real.c:168)
      double *dst = reinterpret_cast<double *>(dstBytes);
      const double *src = static_cast<const double *>(d.val);
      for (int r = 0; r < tl.rows; ++r) dst[r] = src[pr.row0 + r];
      continue;
    }
    T *dst = reinterpret_cast<T *>(dstBytes);
    const T *src = static_cast<const T *>(d.val);
    const int rows = tl.rows, kc = tl.kc;
    for (int r = 0; r < rows; ++r)
      for (int c = 0; c < kc; ++c) {
        const int64_t p = pr.row0 + r, q = pr.col0 + c;
        const int64_t sr = d.swapRC ? q : p, sc = d.swapRC ? p : q;
        dst[(size_t)r * kc + c] = conjIf(src[sr * d.ld + sc], d.conj != 0);
      }
  }
}

void packPool(const Direction &dir, size_t s0, size_t s1, void *dst) {
  for (size_t s = s0; s < s1; ++s) {
    const Stage &S = dir.stages[s];
    char *base = static_cast<char *>(dst) + 16 * (S.poolBegin16 - dir.stages[s0].poolBegin16);
    if (dir.dtype == BF_CUDA_COMPLEX_F64) packStage(dir, S, reinterpret_cast<std::complex<double> *>(base), 0, S.tiles.size(), S.poolBegin16);
    else packStage(dir, S, reinterpret_cast<double *>(base), 0, S.tiles.size(), S.poolBegin16);
  }
}

void packTileRange(const Direction &dir, size_t stage, size_t t0, size_t t1, uint64_t base16, void *dst) {
  const Stage &S = dir.stages[stage];
  if (dir.dtype == BF_CUDA_COMPLEX_F64) packStage(dir, S, reinterpret_cast<std::complex<double> *>(dst), t0, t1, base16);
  else packStage(dir, S, reinterpret_cast<double *>(dst), t0, t1, base16);
}

} // namespace bfc
