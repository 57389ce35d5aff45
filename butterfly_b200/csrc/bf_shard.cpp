// bf_shard.cpp — block-row sharding of one flattened operator over several GPUs: WHO applies WHICH leaf.
// Host only, no CUDA.  ("target-tree subtrees assigned to GPUs: GPU g owns the output rows of its
// subtrees ... and the levels producing them"; the reference itself is single-process, so there is no reference
// code for this — the rules below are ours and tests/test_shard_native.py pins them against the Python
// restatement in butterfly_b200/rowshard.py.)
//
// Every rank keeps the SAME structure (all leaves: identical stages and row addresses everywhere) but applies —
// and stores the factor values of — only the leaves it owns.  Ownership is by contiguous row ranges:
//   * Y is cut into `world` ranges of (nearly) equal rows, snapped to rows no dense leaf straddles;
//   * a MatProduct chain (src/mat_product.c:211-245: its factors and the intermediates between them) whose Y
//     rows lie inside one range lives entirely on that rank — no exchange at all (every near-field leaf and most
//     butterflies of the HODBF, src/fac_helm2.c:806-1002);
//   * a chain whose Y rows span several ranks has every intermediate buffer cut in the same proportions.  Early
//     buffers are ordered by source box, late ones by target box (src/fac_helm2.c:222-401), so a rank's rows of
//     level l mostly depend on its own rows of level l - 1; what does not is pushed to the reader by the
//     producing stage kernel (bf_kernels.cu, peer stores).
// Leaves taller than kMaxLeafRows and the identity / diagonal terms that write Y are cut into row slabs first so
// that the boundaries have somewhere to fall; the summation order of every row is unchanged.
#include "bf_shard.hpp"

#include <algorithm>
#include <cmath>
#include <numeric>

namespace bfc {

namespace {
constexpr int64_t kMaxLeafRows = 2048;

bool storedTransposed(const BfCudaRawOp &o, bool cplx) {
  const uint32_t mask = cplx ? (BF_CUDA_LEAF_TRANS | BF_CUDA_LEAF_CONJ) : BF_CUDA_LEAF_TRANS;
  return (o.flags & mask) != 0;
}

// Slabs of one raw op cut at the acting-row offsets in `bounds` (relative, ascending; those outside (0, m) are
// ignored).  Row-major storage: a slab of acting rows [r, r + h) starts r * ld elements into the stored matrix,
// or r elements for a TRANS-stored leaf (acting rows are stored columns); a diagonal slab starts r doubles in.
// Identity / diagonal leaves act on min(m, k) rows.
void splitOpRows(const BfCudaRawOp &o, const std::vector<int64_t> &bounds, bool cplx, int64_t srcIndex, int64_t srcRow,
                 std::vector<BfCudaRawOp> &out, std::vector<int64_t> &originOp, std::vector<int64_t> &originRow) {
  const int e = cplx ? 2 : 1;
  const bool st = storedTransposed(o, cplx);
  std::vector<int64_t> edges{0};
  for (int64_t b : bounds)
    if (b > 0 && b < o.m && b > edges.back()) edges.push_back(b);
  edges.push_back(o.m);
  for (size_t i = 0; i + 1 < edges.size(); ++i) {
    const int64_t r = edges[i], h = edges[i + 1] - r;
    BfCudaRawOp c = o;
    c.outOff = o.outOff + r;
    c.m = h;
    if (o.kind != BF_CUDA_LEAF_DENSE) {
      if (r >= o.k) break;
      c.inOff = o.inOff + r;
      c.k = std::min(h, o.k - r);
      c.m = c.k;
      if (o.kind == BF_CUDA_LEAF_DIAG && o.val) c.val = static_cast<const double *>(o.val) + r;
    } else if (o.val) {
      c.val = static_cast<const double *>(o.val) + (st ? r * e : r * o.ld * e);
    }
    out.push_back(c);
    originOp.push_back(srcIndex);
    originRow.push_back(srcRow + r);
  }
}

// Rows where a buffer can be cut without splitting a leaf: sorted unique starts s such that every leaf that
// starts earlier ends at or before s.
std::vector<int64_t> cleanCuts(const std::vector<int64_t> &starts, const std::vector<int64_t> &ends) {
  std::vector<size_t> order(starts.size());
  std::iota(order.begin(), order.end(), size_t(0));
  std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return starts[a] < starts[b]; });
  std::vector<int64_t> cuts;
  int64_t run = 0;
  for (size_t i = 0; i < order.size(); ++i) {
    const int64_t s = starts[order[i]];
    if (s >= run && (cuts.empty() || cuts.back() != s)) cuts.push_back(s);
    run = std::max(run, ends[order[i]]);
  }
  return cuts;
}

// Cut [0, rowsTotal) at the clean cuts closest to the cumulative `fractions`.  Returns parts + 1 nondecreasing
// boundaries.
std::vector<int64_t> cutRows(const std::vector<int64_t> &starts, const std::vector<int64_t> &ends, int64_t rowsTotal,
                             const std::vector<double> &fractions) {
  std::vector<int64_t> cand = cleanCuts(starts, ends);
  cand.push_back(rowsTotal);
  double sum = 0;
  for (double f : fractions) sum += f;
  std::vector<int64_t> b{0};
  double cum = 0;
  for (size_t p = 0; p + 1 < fractions.size(); ++p) {
    cum += fractions[p];
    const double target = (sum > 0 ? cum / sum : double(p + 1) / double(fractions.size())) * (double)rowsTotal;
    size_t best = 0;
    double bestD = std::fabs((double)cand[0] - target);
    for (size_t j = 1; j < cand.size(); ++j) {
      const double d = std::fabs((double)cand[j] - target);
      if (d < bestD) { bestD = d; best = j; }
    }
    b.push_back(std::max(cand[best], b.back()));
  }
  b.push_back(rowsTotal);
  return b;
}

int rowOwner(const std::vector<int64_t> &cuts, int64_t r) {
  const int parts = (int)cuts.size() - 1;
  const int i = (int)(std::upper_bound(cuts.begin(), cuts.end(), r) - cuts.begin()) - 1;
  return std::max(0, std::min(i, parts - 1));
}

struct UnionFind {
  std::vector<int64_t> p;
  explicit UnionFind(size_t n) : p(n) { std::iota(p.begin(), p.end(), int64_t(0)); }
  int64_t find(int64_t x) {
    while (p[(size_t)x] != x) { p[(size_t)x] = p[(size_t)p[(size_t)x]]; x = p[(size_t)x]; }
    return x;
  }
  void unite(int64_t a, int64_t b) { a = find(a); b = find(b); if (a != b) p[(size_t)std::max(a, b)] = std::min(a, b); }
};
} // namespace

BfCudaStatus shardOps(const BfCudaPlanDesc &desc, int world, const double *fractionsIn, ShardedOps &S, std::string &err) {
  if (world < 1 || world > kMaxShardRanks) { err = "world size out of range"; return BF_CUDA_ERROR_INVALID_ARGUMENT; }
  if (desc.numBufs < 2 || !desc.bufRows || (desc.numOps > 0 && !desc.ops)) { err = "malformed descriptor"; return BF_CUDA_ERROR_INVALID_ARGUMENT; }
  const bool cplx = desc.dtype == BF_CUDA_COMPLEX_F64;
  std::vector<double> fractions((size_t)world, 1.0);
  if (fractionsIn)
    for (int p = 0; p < world; ++p) fractions[(size_t)p] = fractionsIn[p] > 0 ? fractionsIn[p] : 0.0;
  S = ShardedOps();
  S.world = world;

  // 1. tall dense leaves -> slabs
  std::vector<BfCudaRawOp> ops;
  std::vector<int64_t> oOp, oRow;
  ops.reserve((size_t)desc.numOps);
  for (int64_t i = 0; i < desc.numOps; ++i) {
    const BfCudaRawOp &o = desc.ops[i];
    if (o.flags & BF_CUDA_LEAF_REMOTE) { err = "descriptor is already sharded"; return BF_CUDA_ERROR_INVALID_ARGUMENT; }
    if (o.inBuf < 0 || o.inBuf >= desc.numBufs || o.outBuf < 0 || o.outBuf >= desc.numBufs || o.m < 0 || o.k < 0) {
      err = "op " + std::to_string(i) + ": malformed"; return BF_CUDA_ERROR_INVALID_ARGUMENT;
    }
    std::vector<int64_t> bounds;
    if (o.kind == BF_CUDA_LEAF_DENSE && o.m > kMaxLeafRows)
      for (int64_t r = kMaxLeafRows; r < o.m; r += kMaxLeafRows) bounds.push_back(r);
    splitOpRows(o, bounds, cplx, i, 0, ops, oOp, oRow);
  }

  // 2. Y boundaries: equal shares (or `fractions`) snapped to rows no DENSE leaf writing Y straddles
  {
    std::vector<int64_t> st, en;
    for (const BfCudaRawOp &o : ops)
      if (o.outBuf == BF_CUDA_BUF_Y && o.kind == BF_CUDA_LEAF_DENSE) { st.push_back(o.outOff); en.push_back(o.outOff + o.m); }
    S.yCuts = cutRows(st, en, desc.numRows, fractions);
  }

  // 3. identity / diagonal leaves that write Y (the "I / 2" term of a second-kind system,
  //    multiple_scattering_context.c:483-487) can be cut anywhere: cut them at the Y boundaries
  {
    std::vector<BfCudaRawOp> ops2;
    std::vector<int64_t> oOp2, oRow2;
    ops2.reserve(ops.size());
    for (size_t i = 0; i < ops.size(); ++i) {
      const BfCudaRawOp &o = ops[i];
      std::vector<int64_t> bounds;
      if (o.kind != BF_CUDA_LEAF_DENSE && o.outBuf == BF_CUDA_BUF_Y)
        for (int64_t c : S.yCuts) bounds.push_back(c - o.outOff);
      splitOpRows(o, bounds, cplx, oOp[i], oRow[i], ops2, oOp2, oRow2);
    }
    ops.swap(ops2); oOp.swap(oOp2); oRow.swap(oRow2);
  }

  // 4. owners
  const size_t nOps = ops.size();
  const int64_t nB = desc.numBufs;
  std::vector<int32_t> owner(nOps, -1);
  UnionFind uf((size_t)nB);
  for (const BfCudaRawOp &o : ops)
    if (o.inBuf >= 2 && o.outBuf >= 2) uf.unite(o.inBuf, o.outBuf);
  // chain of an op = component of the temp it touches; ops that read X and write Y directly have none
  std::vector<int64_t> chain(nOps, -1);
  for (size_t i = 0; i < nOps; ++i) {
    const BfCudaRawOp &o = ops[i];
    const int64_t temp = o.outBuf >= 2 ? o.outBuf : (o.inBuf >= 2 ? o.inBuf : -1);
    if (temp >= 0) chain[i] = uf.find(temp);
    else owner[i] = rowOwner(S.yCuts, o.outOff);
  }
  std::vector<size_t> order;
  for (size_t i = 0; i < nOps; ++i) if (chain[i] >= 0) order.push_back(i);
  std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return chain[a] < chain[b]; });
  for (size_t g = 0; g < order.size();) {
    size_t h = g;
    while (h < order.size() && chain[order[h]] == chain[order[g]]) ++h;
    int64_t lo = INT64_MAX, hi = -1;
    for (size_t q = g; q < h; ++q) {
      const BfCudaRawOp &o = ops[order[q]];
      if (o.outBuf == BF_CUDA_BUF_Y) { lo = std::min(lo, o.outOff); hi = std::max(hi, o.outOff + o.m); }
    }
    int p0 = 0, p1 = 0;
    if (hi >= 0) { p0 = rowOwner(S.yCuts, lo); p1 = rowOwner(S.yCuts, std::max(hi - 1, lo)); }
    if (p0 == p1) {
      for (size_t q = g; q < h; ++q) owner[order[q]] = p0;
    } else {
      // the chain's Y rows span ranks p0..p1: cut every intermediate buffer in the proportions of its Y rows
      std::vector<double> frac;
      for (int p = p0; p <= p1; ++p)
        frac.push_back((double)std::max<int64_t>(std::min(S.yCuts[(size_t)p + 1], hi) - std::max(S.yCuts[(size_t)p], lo), 0));
      std::vector<size_t> temps; // ops of the chain that write a temp, grouped by buffer below
      for (size_t q = g; q < h; ++q) {
        const BfCudaRawOp &o = ops[order[q]];
        if (o.outBuf == BF_CUDA_BUF_Y) owner[order[q]] = rowOwner(S.yCuts, o.outOff);
        else temps.push_back(order[q]);
      }
      std::stable_sort(temps.begin(), temps.end(), [&](size_t a, size_t b) { return ops[a].outBuf < ops[b].outBuf; });
      for (size_t t0 = 0; t0 < temps.size();) {
        size_t t1 = t0;
        std::vector<int64_t> st, en;
        while (t1 < temps.size() && ops[temps[t1]].outBuf == ops[temps[t0]].outBuf) {
          st.push_back(ops[temps[t1]].outOff);
          en.push_back(ops[temps[t1]].outOff + ops[temps[t1]].m);
          ++t1;
        }
        const std::vector<int64_t> bc = cutRows(st, en, desc.bufRows[ops[temps[t0]].outBuf], frac);
        for (size_t q = t0; q < t1; ++q) owner[temps[q]] = p0 + rowOwner(bc, ops[temps[q]].outOff);
        t0 = t1;
      }
    }
    g = h;
  }
  for (size_t i = 0; i < nOps; ++i)
    if (owner[i] < 0 || owner[i] >= world) { err = "internal: leaf without an owner"; return BF_CUDA_ERROR_RUNTIME; }
  S.ops.swap(ops);
  S.owner.swap(owner);
  S.originOp.swap(oOp);
  S.originRow.swap(oRow);
  return BF_CUDA_OK;
}

// One rank's view: same leaves, those of other ranks flagged REMOTE; every leaf carries its owner.
void maskForRank(const ShardedOps &S, int rank, std::vector<BfCudaRawOp> &out) {
  out = S.ops;
  for (size_t i = 0; i < out.size(); ++i) {
    out[i].flags &= ~(BF_CUDA_LEAF_REMOTE | (0xFFu << BF_CUDA_LEAF_OWNER_SHIFT));
    out[i].flags |= (uint32_t)S.owner[i] << BF_CUDA_LEAF_OWNER_SHIFT;
    if (S.owner[i] != rank) {
      out[i].flags |= BF_CUDA_LEAF_REMOTE;
      out[i].val = nullptr; // another rank stores these values
    }
  }
}

} // namespace bfc
