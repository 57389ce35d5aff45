// bf_kernels.cu — hand-written sm_100a kernels for the butterfly apply.
//
// Three stage kernels consume the same stage-major tables and the same packed value pool
// (bf_plan.hpp):
//   bfGemvStage     nrhs == 1.  HBM-bandwidth bound (0.5 flop/B complex, 0.25 real).  One warp owns
//                   one static work queue (12 warps / SM) and streams its contiguous span of the pool
//                   through a private 2-slot ring of shared memory filled by TMA bulk copies
//                   (cp.async.bulk + mbarrier complete_tx); the operand segment of every tile rides on
//                   the same mbarrier.  8 (complex) / 16 (real) lanes cooperate on a row: conflict-free
//                   128-byte LDS phases, operand held in registers, one shuffle reduction per ITEM (not
//                   per tile), one coalesced store.  Descriptors live in registers across the lanes.
//                   No atomics: summation order is the plan's order.
//   bfWideMmaStage  complex, nrhs >= 2.  fp64 tensor pipe (DMMA m8n8k4) bound.  Warp-specialised: a
//                   producer warp TMA-stages tiles AND operand rows into one byte ring, four consumer
//                   warps accumulate real / imaginary fragments.
//   bfWideStage     real, nrhs >= 2.  Register-blocked DFMA (the reference has no such apply).
//
// Replaces the per-leaf cblas_zgemm / cblas_dgemv calls of the reference
// (src/mat_dense_complex.c:1754, src/mat_dense_real.c:1358,1428) and the block loops around
// them (src/mat_block_coo.c:404-418, src/mat_block_diag.c:387-399, src/mat_block_dense.c:534-566).
#include "bf_kernels.cuh"

namespace bfc {

// ---- PTX helpers ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smemU32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbarExpectTx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ uint64_t policyEvictFirst() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// 1-D TMA bulk copy global -> shared, completion counted in bytes on an mbarrier.
__device__ __forceinline__ void bulkLoad(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulkLoadHint(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, uint64_t pol) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(pol) : "memory");
}
__device__ __forceinline__ void fenceProxyAsyncGlobal() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ uint32_t loadAcquire(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fenceBarrierInit() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fenceProxyAsync() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ uint32_t tileBytes(int es, uint32_t rows, uint32_t kc) {
  if (kc == kKcIdentity) return 0u;
  if (kc == kKcDiag) return (rows * 8u + 15u) & ~15u;
  return (rows * kc * (uint32_t)es + 15u) & ~15u;
}

template <bool CPLX> struct Scalar;
template <> struct Scalar<true> {
  using T = double2;
  static __device__ __forceinline__ T zero() { return make_double2(0.0, 0.0); }
  static __device__ __forceinline__ void fma(T &acc, const T a, const T x) {
    acc.x = ::fma(a.x, x.x, acc.x); acc.x = ::fma(-a.y, x.y, acc.x);
    acc.y = ::fma(a.x, x.y, acc.y); acc.y = ::fma(a.y, x.x, acc.y);
  }
  static __device__ __forceinline__ void fmaReal(T &acc, double d, const T x) { acc.x = ::fma(d, x.x, acc.x); acc.y = ::fma(d, x.y, acc.y); }
  static __device__ __forceinline__ T add(T a, T b) { return make_double2(a.x + b.x, a.y + b.y); }
  static __device__ __forceinline__ T shflXor(T v, int m) { return make_double2(__shfl_xor_sync(0xffffffffu, v.x, m), __shfl_xor_sync(0xffffffffu, v.y, m)); }
  static __device__ __forceinline__ T shfl(T v, int src) { return make_double2(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src)); }
};
template <> struct Scalar<false> {
  using T = double;
  static __device__ __forceinline__ T zero() { return 0.0; }
  static __device__ __forceinline__ void fma(T &acc, const T a, const T x) { acc = ::fma(a, x, acc); }
  static __device__ __forceinline__ void fmaReal(T &acc, double d, const T x) { acc = ::fma(d, x, acc); }
  static __device__ __forceinline__ T add(T a, T b) { return a + b; }
  static __device__ __forceinline__ T shflXor(T v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
  static __device__ __forceinline__ T shfl(T v, int src) { return __shfl_sync(0xffffffffu, v, src); }
};

__device__ __forceinline__ unsigned char *rowPtr(const StageLaunch &a, uint32_t addr, int es) {
  const uint32_t sp = addr >> 30;
  return a.base[sp] + (long long)(addr & kRowMask) * a.ld[sp] * es;
}

// acc[p] += A[row(p), :] . x for every pass p of one staged tile.  NC = ceil(kc / G) register
// chunks.  Lanes past the tile edge read a clamped (valid) element and multiply it by a zero
// operand (column edge) or feed an accumulator nobody stores (row edge), so the loop body has no
// predicates; passes are taken two at a time for ILP and left through a warp-uniform branch.
template <bool CPLX, int NC, int P, int RPP, int G>
__device__ __forceinline__ void tileFma(typename Scalar<CPLX>::T (&acc)[P], const typename Scalar<CPLX>::T *A,
                                        const typename Scalar<CPLX>::T *xs, const int (&rp)[P], int rows, int kc, int l) {
  using S = Scalar<CPLX>;
  using T = typename S::T;
  // Only the LAST column chunk can run past the tile edge (NC = ceil(kc / G)): the others are addressed as
  // base + immediate, one address register per pass instead of one per load.  rp[p] is the lane's (clamped)
  // row of pass p, fixed for the whole item.
  T xr[NC];
  const int jl = (NC - 1) * G + l;
  const int cl = min(jl, kc - 1);
#pragma unroll
  for (int tt = 0; tt < NC - 1; ++tt) xr[tt] = xs[tt * G + l];
  {
    const T v = xs[cl];
    xr[NC - 1] = jl < kc ? v : S::zero();
  }
  const T *Al = A + l;
#pragma unroll
  for (int p = 0; p < P; p += 2) {
    if (p * RPP >= rows) break;
    const int o0 = rp[p] * kc, o1 = rp[p + 1] * kc;
    T a0[NC], a1[NC];
#pragma unroll
    for (int tt = 0; tt < NC - 1; ++tt) { a0[tt] = Al[o0 + tt * G]; a1[tt] = Al[o1 + tt * G]; }
    a0[NC - 1] = A[o0 + cl];
    a1[NC - 1] = A[o1 + cl];
#pragma unroll
    for (int tt = 0; tt < NC; ++tt) { S::fma(acc[p], a0[tt], xr[tt]); S::fma(acc[p + 1], a1[tt], xr[tt]); }
  }
}

// ---- 1-RHS stage kernel ---------------------------------------------------------------------------
//
// FUSED = the whole apply in ONE launch (bf_plan.hpp, FusedTables): the tables are queue-major over all
// stages, a warp never leaves its ring, and the grid-wide barrier between stages is replaced by readiness
// counters — lane 0 checks (ld.acquire) that every writer of a tile's operand buffer has signalled before it
// issues the tile's bulk copies; an item signals (fence + atomic add) after its store.  A tile whose operand is
// not ready yet is simply not staged (the warp goes on consuming what it has) unless the consumer has nothing
// else to do, in which case it spins: all CTAs are co-resident (cooperative launch, one CTA per SM) and every
// wait is on items of an EARLIER stage, which a warp reaches without waiting on anything of this one, so the
// spin always ends.  The counters of the NEXT apply are zeroed here (two sets, alternating).
//
// SHARDED (MODE 2) = one stage of one rank of a block-row sharded operator (ShardLaunch): before it touches the
// workspace every warp waits for the flags of the ranks that pushed rows to this one during the previous stage
// (stage 0: for every peer to have finished the previous apply — the workspace is about to be overwritten);
// an item whose rows another rank reads stores them into that rank's workspace as well; the last warp of the
// grid to finish posts this stage's flag to every peer.
template <bool CPLX, int MODE>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 1) bfGemvStage(const __grid_constant__ StageLaunch a) {
  constexpr bool FUSED = MODE == 1, SHARDED = MODE == 2;
  using S = Scalar<CPLX>;
  using T = typename S::T;
  constexpr int ES = CPLX ? 16 : 8;
  constexpr int G = CPLX ? 8 : 16; // lanes per row: one 128-byte LDS phase per row chunk
  constexpr int RPP = 32 / G;      // rows per pass
  constexpr int P = 8;             // passes: RPP * P == max item rows
  constexpr int KT = 8;            // register chunks of G columns: G * KT == max tile columns
  static_assert(RPP * P == (CPLX ? kMaxRowsComplex : kMaxRowsReal), "item rows");
  static_assert(G * KT >= (CPLX ? kMaxKcComplex : kMaxKcReal), "tile columns");
  static_assert((CPLX ? kMaxKcComplex * 16 : kMaxKcReal * 8 + 16) <= kGemvXBytes, "x slot");

  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char *ring = smem + warp * (kGemvSlots * kGemvSlotBytes);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kWarpsPerCta * kGemvSlots * kGemvSlotBytes) + warp * kGemvSlots;
  // shared-window addresses once (a generic -> shared conversion per use cost ~13 instructions per tile)
  const uint32_t ringU = smemU32(ring), barsU = smemU32(bars);
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kGemvSlots; ++s) mbarInit(barsU + 8u * s, 1);
    fenceBarrierInit();
  }
  if (SHARDED && threadIdx.x == 0) {
    // Block-row shard: the rows peers pushed during the previous stage have landed once their flags show this
    // apply's epoch (a flag is written after the rows).  Stage 0 instead waits for EVERY peer to have finished the
    // previous apply (their last stage's flag, epoch - 1): its workspace rows are about to be overwritten.
    // ONE thread per CTA polls, with relaxed loads and a single fence at the end; the CTA barrier below hands the
    // result to the other warps.  (Every warp polling with ld.acquire.sys — 1 776 pollers on one line — cost
    // 0.4 - 1 ms per apply at N = 524 288.)  Every CTA waits, also one without work: the NEXT stage relies on this
    // grid having seen the previous pushes.
    const ShardLaunch &sh = *a.shard;
    const int ws = a.stage == 0 ? sh.numStages - 1 : a.stage - 1;
    const uint32_t want = a.stage == 0 ? a.epoch - 1u : a.epoch;
    uint32_t mask = a.stage == 0 ? ((1u << sh.world) - 1u) & ~(1u << sh.rank) : sh.recvMask[ws];
    const volatile uint32_t *flags = sh.peerFlags[sh.rank] + (size_t)ws * sh.world;
    for (; mask; mask &= mask - 1) {
      const int r = __ffs(mask) - 1;
      for (unsigned long long spins = 0; (int32_t)(flags[r] - want) < 0; ++spins) {
        if (spins > (1ull << 25)) __trap(); // a peer that never arrives: fail loudly instead of hanging
        __nanosleep(100);
      }
    }
    __threadfence_system(); // acquire: the rows behind the flags
  }
  __syncthreads();
  if (SHARDED && lane == 0) fenceProxyAsyncGlobal(); // peer stores (generic proxy) -> this lane's bulk copies

  // Under programmatic dependent launch EVERY exit path runs griddepcontrol.wait first: stage-to-stage
  // ordering is transitive only through grids whose threads all waited, and a stage with no local items
  // (a row-sharded plan whose leaves of that stage are all remote) must not let stage s + 1 overtake s - 1.
  auto pdlSync = [&]() {
    if (a.pdl) {
      asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
      asm volatile("griddepcontrol.wait;" ::: "memory");
    }
  };
  const int q = blockIdx.x * kWarpsPerCta + warp;
  if (FUSED) { // zero the counter set of the next apply (the previous user of that set has completed: stream order)
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.numCounters; i += gridDim.x * blockDim.x) a.cntNext[i] = 0u;
  }
  // block-row shard: hand-over between ranks
  bool pushedToPeers = false; // this warp stored rows into another rank's workspace
  auto shardDone = [&]() { // this warp's stores (local and to peers) are complete; the last warp posts the flags
    if (!SHARDED) return;
    __syncwarp();
    if (lane == 0) {
      const ShardLaunch &sh = *a.shard;
      // Only a warp that stored into a peer needs its stores performed system-wide before it counts itself done
      // (rows kept on this GPU are handed to the next stage by the kernel boundary).  A system fence in every warp
      // — 1 776 of them arriving together — cost ~20 us per stage.
      if (pushedToPeers) __threadfence_system();
      else __threadfence();
      const uint32_t old = atomicAdd(sh.done + a.stage, 1u);
      if ((old + 1u) % sh.totalWarps == 0u) {
        // ONE system fence (release: every warp's rows, observed through the counter above), then plain flag
        // stores — to the ranks that wait on this stage only; after the last stage to everybody (stage 0 of the
        // next apply waits for all peers).  Seven st.release.sys in a row cost ~15 us per stage.
        __threadfence_system();
        uint32_t to = a.stage == sh.numStages - 1 ? ((1u << sh.world) - 1u) & ~(1u << sh.rank) : sh.sendMask[a.stage];
        for (; to; to &= to - 1) {
          const int r = __ffs(to) - 1;
          uint32_t *f = sh.peerFlags[r] + (size_t)a.stage * sh.world + sh.rank;
          asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(f), "r"(a.epoch) : "memory");
        }
      }
    }
  };
  if (q >= a.numQueues) { pdlSync(); shardDone(); return; }
  const uint32_t iBeg = a.qItemBegin[q], iEnd = a.qItemBegin[q + 1];
  const uint32_t tBeg = a.qTileBegin[q], tEnd = a.qTileBegin[q + 1];
  if (iBeg == iEnd) { pdlSync(); shardDone(); return; }
  const uint64_t pol = policyEvictFirst();

  // Descriptor windows: lane l holds tiles[twBase + l] / tiles[twBase + 32 + l] (and the same for
  // items), refilled 32 entries ahead of use, so that neither cursor ever waits on a table load —
  // a descriptor is one shuffle away.
  const uint64_t *tileTab = reinterpret_cast<const uint64_t *>(a.tiles);
  const uint4 *itemTab = reinterpret_cast<const uint4 *>(a.items);
  auto loadTiles = [&](uint32_t base) -> uint64_t { return base + lane < tEnd ? tileTab[base + lane] : 0ull; };
  auto loadItems = [&](uint32_t base) -> uint4 { return base + lane < iEnd ? itemTab[base + lane] : make_uint4(0u, 0u, 0u, 0u); };
  // fused tables: the dependency words travel in windows of their own, slid together with the descriptors
  const uint64_t *tileDepTab = reinterpret_cast<const uint64_t *>(a.tileDeps);
  const uint64_t *itemDepTab = reinterpret_cast<const uint64_t *>(a.itemDeps);
  auto loadTileDeps = [&](uint32_t base) -> uint64_t { return (FUSED && base + lane < tEnd) ? tileDepTab[base + lane] : ~0ull; };
  auto loadItemDeps = [&](uint32_t base) -> uint64_t {
    if (FUSED) return base + lane < iEnd ? itemDepTab[base + lane] : ~0ull;
    if (SHARDED) return base + lane < iEnd ? (uint64_t)a.itemPeers[base + lane] : 0ull; // ranks that read the item's rows
    return ~0ull;
  };
  uint32_t twBase = tBeg, iwBase = iBeg;
  uint64_t tw0 = loadTiles(twBase), tw1 = loadTiles(twBase + 32);
  uint4 iw0 = loadItems(iwBase), iw1 = loadItems(iwBase + 32);
  uint64_t td0 = loadTileDeps(twBase), td1 = loadTileDeps(twBase + 32);
  uint64_t id0 = loadItemDeps(iwBase), id1 = loadItemDeps(iwBase + 32);
  auto getTile = [&](uint32_t t) -> DevTile {
    const uint32_t off = t - twBase; // warp-uniform
    uint64_t v;
    if (off < 32) v = __shfl_sync(0xffffffffu, tw0, off);
    else if (off < 64) v = __shfl_sync(0xffffffffu, tw1, off - 32);
    else v = tileTab[t]; // the producer ran far ahead over value-less tiles
    DevTile d;
    d.inAddr = (uint32_t)v; d.kc = (uint16_t)(v >> 32); d.rows = (uint16_t)(v >> 48);
    return d;
  };
  auto getTileDep = [&](uint32_t t) -> uint64_t { // low word: counter, high word: its final value
    const uint32_t off = t - twBase;
    if (off < 32) return __shfl_sync(0xffffffffu, td0, off);
    if (off < 64) return __shfl_sync(0xffffffffu, td1, off - 32);
    return tileDepTab[t];
  };
  uint64_t curItemDep = ~0ull;
  auto getItem = [&](uint32_t i) -> DevItem {
    if (i - iwBase >= 32) {
      iw0 = iw1; iwBase += 32; iw1 = loadItems(iwBase + 32);
      if (FUSED || SHARDED) { id0 = id1; id1 = loadItemDeps(iwBase + 32); }
    }
    const uint32_t off = i - iwBase;
    if (FUSED || SHARDED) curItemDep = __shfl_sync(0xffffffffu, id0, off);
    DevItem d;
    d.outAddr = __shfl_sync(0xffffffffu, iw0.x, off);
    d.tileBegin = __shfl_sync(0xffffffffu, iw0.y, off);
    const uint32_t rf = __shfl_sync(0xffffffffu, iw0.z, off);
    d.rows = (uint16_t)rf; d.flags = (uint16_t)(rf >> 16);
    d.numTiles = __shfl_sync(0xffffffffu, iw0.w, off);
    return d;
  };

  // producer (lane 0 issues; every lane tracks the same scalars) and consumer cursors
  uint32_t pT = tBeg, pSeq = 0, cSeq = 0;
  // (fused tables: every span of the queue opens with a jump tile that sets both pool cursors)
  const unsigned char *pPool = FUSED ? a.pool : a.pool + a.qPool16[q] * 16ull;
  const unsigned char *cPool = pPool;

  // operand rows of the 1-RHS kernel are dense (ld == 1: bf_capi.cu stages X otherwise, the workspace
  // has one column), so a row address is base + row * ES with the base picked from two registers
  const unsigned char *const baseX = a.base[kSpaceX], *const baseW = a.base[kSpaceW];
  auto inRowPtr = [&](uint32_t addr) -> const unsigned char * {
    return ((addr >> 30) == kSpaceX ? baseX : baseW) + (size_t)(addr & kRowMask) * ES;
  };
  // readiness counters (FUSED).  Signals are DEFERRED and coalesced: an item's bump is posted when a later item
  // signals another counter (consecutive items mostly write the same buffer: one atomic for the run), when the
  // warp is about to spin, or at the end of the queue.  By then the item's stores have long been acknowledged,
  // so the fence ahead of the atomic does not stall the warp the way a fence right behind the stores does
  // (measured: signalling every item at once cost 19 % at N = 524 288).  Posting before every spin keeps the
  // scheme deadlock-free: a warp never waits while it sits on a signal somebody else may be waiting for.
  uint32_t pendIdx = kNoCounter, pendCount = 0;
  auto flushSignal = [&]() {
    if (pendIdx != kNoCounter) {
      if (lane == 0) {
        __threadfence();
        atomicAdd(a.cnt + pendIdx, pendCount);
      }
      pendIdx = kNoCounter;
      pendCount = 0;
    }
  };
  // spin = wait until counter `idx` has reached `target`; otherwise one look
  uint32_t readyIdx = kNoCounter; // the last operand-buffer counter seen complete (tiles of an item mostly share it)
  auto counterReached = [&](uint32_t idx, uint32_t target, bool spin) -> bool {
    uint32_t v = 0;
    if (spin) flushSignal();
    if (lane == 0) {
      v = loadAcquire(a.cnt + idx);
      // watchdog: a wait that outlasts any real apply (seconds) is a broken plan, not a slow producer —
      // fail the launch loudly instead of hanging the device
      for (unsigned long long spins = 0; spin && v < target; ++spins) {
        if (spins > (1ull << 25)) __trap();
        __nanosleep(40);
        v = loadAcquire(a.cnt + idx);
      }
    }
    v = __shfl_sync(0xffffffffu, v, 0);
    return v >= target;
  };
  // Stage the next tile of the queue.  Returns false when there is none left, or (FUSED, !block) when its
  // operand is not complete yet — the cursor then stays on it.
  auto produce = [&](bool block) -> bool {
    while (pT < tEnd) {
      const DevTile t = getTile(pT);
      if (t.kc >= kKcJump) { // identity / diagonal tiles are read directly by the consumer; jumps move the cursor
        if (t.kc == kKcJump) pPool = a.pool + (((unsigned long long)t.rows << 32) | t.inAddr) * 16ull;
        else pPool += tileBytes(ES, t.rows, t.kc);
        ++pT;
        continue;
      }
      if (FUSED) {
        const uint64_t dep = getTileDep(pT);
        const uint32_t idx = (uint32_t)dep;
        if (idx != kNoCounter && idx != readyIdx) {
          if (!counterReached(idx, (uint32_t)(dep >> 32), block)) return false;
          readyIdx = idx;
          if (lane == 0) fenceProxyAsyncGlobal(); // generic-proxy stores of other SMs -> this thread's bulk copies
        }
      }
      ++pT;
      const uint32_t bytes = CPLX ? (uint32_t)t.rows * t.kc * 16u : ((uint32_t)t.rows * t.kc * 8u + 15u) & ~15u;
      const unsigned char *src = pPool;
      pPool += bytes;
      if (lane == 0) {
        const uint32_t slot = pSeq % kGemvSlots;
        const uint32_t dst = ringU + slot * kGemvSlotBytes;
        const uint32_t bar = barsU + 8u * slot;
        const unsigned char *xsrc = inRowPtr(t.inAddr);
        uint32_t xb;
        if (CPLX) {
          xb = (uint32_t)t.kc * 16u;
        } else { // 8-byte elements: copy from the enclosing 16-byte aligned window
          const uint32_t shift = (uint32_t)(reinterpret_cast<uintptr_t>(xsrc) & 8u);
          xsrc -= shift;
          xb = (shift + (uint32_t)t.kc * 8u + 15u) & ~15u;
        }
        mbarExpectTx(bar, bytes + xb);
        bulkLoadHint(dst, src, bytes, bar, pol);
        bulkLoad(dst + kTileBytes, xsrc, xb, bar);
      }
      ++pSeq;
      return true;
    }
    return false;
  };
  // Programmatic dependent launch (small operators only, bf_capi.cu): everything above (barriers, queue
  // bounds, descriptor windows — static plan tables) may overlap the tail of the previous stage; its results
  // (the operand rows the first bulk copy reads, the rows this stage overwrites) are only touched after the
  // wait.  With one 222 KiB CTA per SM the overlap is the prologue of a CTA on an SM the previous stage has
  // already left.  Measured A/B: 0.175 -> 0.165 ms per apply at N = 16 384, 0.630 -> 0.608 at 65 536, but
  // 10.67 -> 10.85 ms at N = 524 288 — hence the size switch.
  pdlSync();
#pragma unroll 1
  for (int s = 0; s < kGemvSlots; ++s)
    if (!produce(false)) break;

  const int g = lane / G, l = lane % G;
#pragma unroll 1
  for (uint32_t i = iBeg; i < iEnd; ++i) {
    const DevItem it = getItem(i);
    const int rows = it.rows;
    T acc[P];
#pragma unroll
    for (int p = 0; p < P; ++p) acc[p] = S::zero();
    T extra = S::zero();
    int rp[P]; // this lane's row in every pass, clamped to the item's last row (edge lanes repeat it)
#pragma unroll
    for (int p = 0; p < P; ++p) rp[p] = min(p * RPP + g, rows - 1);

#pragma unroll 1
    for (uint32_t t = it.tileBegin; t < it.tileBegin + it.numTiles; ++t) {
      if (t - twBase >= 32 && pT - twBase >= 32) { // both cursors left window 0
        tw0 = tw1; twBase += 32; tw1 = loadTiles(twBase + 32);
        if (FUSED) { td0 = td1; td1 = loadTileDeps(twBase + 32); }
      }
      const DevTile tl = getTile(t);
      if (tl.kc >= kKcJump) {
        if (tl.kc == kKcJump) {
          cPool = a.pool + (((unsigned long long)tl.rows << 32) | tl.inAddr) * 16ull;
          continue;
        }
        if (FUSED) { // read with plain loads: wait for the operand here (L1 is bypassed: other SMs wrote it)
          const uint64_t dep = getTileDep(t);
          if ((uint32_t)dep != kNoCounter && (uint32_t)dep != readyIdx) {
            counterReached((uint32_t)dep, (uint32_t)(dep >> 32), true);
            readyIdx = (uint32_t)dep;
          }
        }
        if (lane < rows) {
          const T *xp = reinterpret_cast<const T *>(inRowPtr(tl.inAddr) + (size_t)lane * ES);
          const T xv = FUSED ? __ldcg(xp) : *xp;
          const double d = tl.kc == kKcDiag ? reinterpret_cast<const double *>(cPool)[lane] : 1.0;
          S::fmaReal(extra, d, xv);
        }
        cPool += tileBytes(ES, tl.rows, tl.kc);
        continue;
      }
      if (FUSED) { // the tile must have been staged: now the consumer has nothing else to do, so spin if need be
        while (pSeq == cSeq) produce(true);
      }
      const int kc = tl.kc;
      const uint32_t bytes = CPLX ? (uint32_t)rows * kc * 16u : ((uint32_t)rows * kc * 8u + 15u) & ~15u;
      const uint32_t slot = cSeq % kGemvSlots;
      const unsigned char *As = ring + slot * kGemvSlotBytes;
      const unsigned char *xsB = As + kTileBytes;
      if (!CPLX) xsB += reinterpret_cast<uintptr_t>(inRowPtr(tl.inAddr)) & 8u;
      mbarWait(barsU + 8u * slot, (cSeq / kGemvSlots) & 1u);

      const T *xs = reinterpret_cast<const T *>(xsB);
      const T *A = reinterpret_cast<const T *>(As);
      // exact trip counts through real (warp-uniform) branches: no predicated-off FMA slots
      switch ((kc + G - 1) / G) {
      case 1: tileFma<CPLX, 1, P, RPP, G>(acc, A, xs, rp, rows, kc, l); break;
      case 2: tileFma<CPLX, 2, P, RPP, G>(acc, A, xs, rp, rows, kc, l); break;
      case 3: tileFma<CPLX, 3, P, RPP, G>(acc, A, xs, rp, rows, kc, l); break;
      case 4: tileFma<CPLX, 4, P, RPP, G>(acc, A, xs, rp, rows, kc, l); break;
      case 5: tileFma<CPLX, 5, P, RPP, G>(acc, A, xs, rp, rows, kc, l); break;
      case 6: tileFma<CPLX, 6, P, RPP, G>(acc, A, xs, rp, rows, kc, l); break;
      case 7: tileFma<CPLX, 7, P, RPP, G>(acc, A, xs, rp, rows, kc, l); break;
      default: tileFma<CPLX, 8, P, RPP, G>(acc, A, xs, rp, rows, kc, l); break;
      }
      __syncwarp();
      ++cSeq;
      cPool += bytes;
      if (FUSED) { // refill what is free; a tile whose operand is not complete yet stays where it is
        while (pSeq - cSeq < (uint32_t)kGemvSlots && produce(false)) {}
      } else {
        produce(false); // refill the slot just drained
      }
    }
    if (FUSED && (it.flags & kItemJump)) continue; // pseudo item: nothing to reduce or store

    // One reduction per item over the G lanes of a row, then row r -> lane r for a coalesced store.  The 8
    // passes are shared out over the lanes as the butterfly halves (lane l keeps the passes whose bit matches
    // its own and hands the others to its partner): 4 + 2 + 1 exchanges instead of 8 x 3, and every sum is
    // formed with the same operand pairs in the same level order as a plain xor all-reduce, i.e. bit-identical.
    static_assert(P == 8, "the halving butterfly below shares out 8 passes");
    if (G == 16) {
#pragma unroll
      for (int p = 0; p < P; ++p) acc[p] = S::add(acc[p], S::shflXor(acc[p], 8));
    }
    T r4[4], r2[2];
    {
      const bool hi = (l & 4) != 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) r4[j] = S::add(hi ? acc[j + 4] : acc[j], S::shflXor(hi ? acc[j] : acc[j + 4], 4));
    }
    {
      const bool hi = (l & 2) != 0;
#pragma unroll
      for (int j = 0; j < 2; ++j) r2[j] = S::add(hi ? r4[j + 2] : r4[j], S::shflXor(hi ? r4[j] : r4[j + 2], 2));
    }
    const bool odd = (l & 1) != 0;
    const T mine = S::add(odd ? r2[1] : r2[0], S::shflXor(odd ? r2[0] : r2[1], 1));
    // lane (g, l) now holds pass (l & 7) of row group g, i.e. row (l & 7) * RPP + g; row r lives in lane
    // (r % RPP) * G + r / RPP
    T out = S::shfl(mine, (lane % RPP) * G + lane / RPP);
    out = S::add(out, extra);
    if (FUSED && (it.flags & kItemAccumulate) && (uint32_t)curItemDep != kNoCounter)
      counterReached((uint32_t)curItemDep, (uint32_t)(curItemDep >> 32), true); // the earlier stages' share of these rows
    if (lane < rows) {
      const uint32_t sp = it.outAddr >> 30;
      long long row = (long long)(it.outAddr & kRowMask) + lane;
      if (sp == kSpaceY && a.outPerm) row = a.outPerm[row];
      T *dst = reinterpret_cast<T *>(a.base[sp] + row * a.ld[sp] * ES);
      if (it.flags & kItemAccumulate) out = S::add(out, FUSED ? __ldcg(dst) : *dst);
      *dst = out;
      if (SHARDED) { // the ranks that read these rows get them straight into their workspace (same row address)
        if ((uint32_t)curItemDep) pushedToPeers = true; // (warp-uniform: the mask belongs to the item)
        for (uint32_t m = (uint32_t)curItemDep; m; m &= m - 1) {
          const int r = __ffs(m) - 1;
          *reinterpret_cast<T *>(a.shard->peerW[r] + row * ES) = out;
        }
      }
    }
    if (FUSED && (uint32_t)curItemDep != kNoCounter) { // signal: the rows are complete as far as this stage goes
      __syncwarp(); // the stores of all lanes are ordered before whatever lane 0 does next
      if (pendIdx != (uint32_t)curItemDep) {
        flushSignal();
        pendIdx = (uint32_t)curItemDep;
      }
      ++pendCount;
    }
  }
  if (FUSED) flushSignal();
  shardDone();
}

// ---- many-RHS stage kernel ------------------------------------------------------------------------
// Register-blocked: a CTA of 4 warps owns an item (<= R rows) x (32 * CT) RHS columns.  Thread
// (ty = warp, tx = lane) accumulates rows ty, ty + 4, ... against columns tx, tx + 32, ...: every A
// element is one warp-uniform LDS broadcast reused for CT columns, every operand element one
// coalesced global load reused for RT rows, i.e. 4 * RT * CT DFMA per (RT + CT) loads per k-step
// for complex.  fp64-pipe bound; the A tiles arrive through the same TMA ring as in
// the 1-RHS kernel, one ring per CTA.
template <bool CPLX, int CT>
__global__ void __launch_bounds__(kWideThreads) bfWideStage(const __grid_constant__ StageLaunch a) {
  using S = Scalar<CPLX>;
  using T = typename S::T;
  constexpr int ES = CPLX ? 16 : 8;
  constexpr int R = CPLX ? kMaxRowsComplex : kMaxRowsReal;
  constexpr int TY = kWideThreads / 32;
  constexpr int RT = R / TY;
  constexpr int NC = 32 * CT;

  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kWideSlots * kTileBytes);
  const int tid = threadIdx.x;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < kWideSlots; ++s) mbarInit(smemU32(&bars[s]), 1);
    fenceBarrierInit();
  }
  __syncthreads();

  const int q0 = blockIdx.x * kWideQueuesPerCta;
  if (q0 >= a.numQueues) return;
  const int q1 = min(q0 + kWideQueuesPerCta, a.numQueues);
  const uint32_t iBeg = a.qItemBegin[q0], iEnd = a.qItemBegin[q1];
  const uint32_t tEnd = a.qTileBegin[q1];
  if (iBeg == iEnd) return;
  const uint64_t pol = policyEvictFirst();

  const int tx = tid & 31, ty = tid >> 5;
  const long long col0 = (long long)a.colBegin + (long long)blockIdx.y * NC + tx;
  bool colOk[CT];
#pragma unroll
  for (int c = 0; c < CT; ++c) colOk[c] = col0 + 32 * c < a.nrhs;

  uint32_t pT = a.qTileBegin[q0], pSeq = 0, cSeq = 0;
  const unsigned char *pPool = a.pool + a.qPool16[q0] * 16ull;
  const unsigned char *cPool = pPool;
  auto produce = [&]() {
    while (pT < tEnd) {
      const DevTile t = a.tiles[pT];
      ++pT;
      const uint32_t bytes = tileBytes(ES, t.rows, t.kc);
      const unsigned char *src = pPool;
      pPool += bytes;
      if (t.kc >= kKcDiag) continue;
      if (tid == 0) {
        const uint32_t slot = pSeq % kWideSlots;
        const uint32_t bar = smemU32(&bars[slot]);
        mbarExpectTx(bar, bytes);
        bulkLoadHint(smemU32(smem + slot * kTileBytes), src, bytes, bar, pol);
      }
      ++pSeq;
      return;
    }
  };
#pragma unroll 1
  for (int s = 0; s < kWideSlots; ++s) produce();

#pragma unroll 1
  for (uint32_t i = iBeg; i < iEnd; ++i) {
    const DevItem it = a.items[i];
    const int rows = it.rows;
    T acc[RT][CT];
#pragma unroll
    for (int u = 0; u < RT; ++u)
#pragma unroll
      for (int c = 0; c < CT; ++c) acc[u][c] = S::zero();

#pragma unroll 1
    for (uint32_t t = it.tileBegin; t < it.tileBegin + it.numTiles; ++t) {
      const DevTile tl = a.tiles[t];
      const uint32_t bytes = tileBytes(ES, tl.rows, tl.kc);
      const uint32_t sp = tl.inAddr >> 30;
      const long long ldIn = a.ld[sp];
      const T *xp = reinterpret_cast<const T *>(a.base[sp]) + (long long)(tl.inAddr & kRowMask) * ldIn + col0;
      if (tl.kc >= kKcDiag) {
        const double *dg = reinterpret_cast<const double *>(cPool);
#pragma unroll
        for (int u = 0; u < RT; ++u) {
          const int r = u * TY + ty;
          if (r < rows) {
            const double d = tl.kc == kKcDiag ? dg[r] : 1.0;
#pragma unroll
            for (int c = 0; c < CT; ++c)
              if (colOk[c]) S::fmaReal(acc[u][c], d, xp[(long long)r * ldIn + 32 * c]);
          }
        }
        cPool += bytes;
        continue;
      }
      const int kc = tl.kc;
      const uint32_t slot = cSeq % kWideSlots;
      const T *A = reinterpret_cast<const T *>(smem + slot * kTileBytes);

      // The operand rows of the NEXT U steps are fetched while these U compute; the sums run over j in order.
      // ncu (profiles/r2zz_realwide_summary.txt): the loop was ISSUE-bound, not latency-bound — fp64 pipe active 20 %,
      // issue slots 62 %, i.e. five other instructions per DFMA (per-load guards, 64-bit index multiplies, a
      // row-count branch per row and step).  Hence: whole blocks of U steps run unguarded on precomputed pointers,
      // only the last block of a tile and items of fewer than R rows take the guarded path.
      constexpr int U = CT <= 2 ? 8 : 4;
      const T *xq[CT]; // next operand row to fetch, per column of this thread
#pragma unroll
      for (int c = 0; c < CT; ++c) xq[c] = xp + 32 * c;
      T xa[U][CT];
      auto fetch = [&](T (&dst)[U][CT], int jb) { // rows jb .. jb + U - 1 of the tile's operand segment
        if (jb + U <= kc) {
#pragma unroll
          for (int s = 0; s < U; ++s)
#pragma unroll
            for (int c = 0; c < CT; ++c) dst[s][c] = colOk[c] ? xq[c][(long long)s * ldIn] : S::zero();
        } else {
#pragma unroll
          for (int s = 0; s < U; ++s)
#pragma unroll
            for (int c = 0; c < CT; ++c) dst[s][c] = (colOk[c] && jb + s < kc) ? xq[c][(long long)s * ldIn] : S::zero();
        }
#pragma unroll
        for (int c = 0; c < CT; ++c) xq[c] += (long long)U * ldIn;
      };
      fetch(xa, 0);
      mbarWait(smemU32(&bars[slot]), (cSeq / kWideSlots) & 1u);
      const T *Ap[RT]; // this thread's rows of the tile; rows past the edge are clamped (never stored)
#pragma unroll
      for (int u = 0; u < RT; ++u) Ap[u] = A + min(u * TY + ty, rows - 1) * kc;
      const bool fullRows = (RT - 1) * TY < rows; // every row group of the thread tile exists (warp-uniform)
#pragma unroll 1
      for (int j0 = 0; j0 < kc; j0 += U) {
        T xn[U][CT];
        fetch(xn, j0 + U);
        if (fullRows && j0 + U <= kc) {
#pragma unroll
          for (int s = 0; s < U; ++s)
#pragma unroll
            for (int u = 0; u < RT; ++u) {
              const T av = Ap[u][j0 + s];
#pragma unroll
              for (int c = 0; c < CT; ++c) S::fma(acc[u][c], av, xa[s][c]);
            }
        } else {
#pragma unroll
          for (int s = 0; s < U; ++s) {
            if (j0 + s < kc) { // warp-uniform
#pragma unroll
              for (int u = 0; u < RT; ++u) {
                if (u * TY < rows) { // warp-uniform
                  const T av = Ap[u][j0 + s];
#pragma unroll
                  for (int c = 0; c < CT; ++c) S::fma(acc[u][c], av, xa[s][c]);
                }
              }
            }
          }
        }
#pragma unroll
        for (int s = 0; s < U; ++s)
#pragma unroll
          for (int c = 0; c < CT; ++c) xa[s][c] = xn[s][c];
      }
      __syncthreads(); // every warp is done with this slot
      produce();
      ++cSeq;
      cPool += bytes;
    }

    const uint32_t sp = it.outAddr >> 30;
    const long long ldOut = a.ld[sp];
#pragma unroll
    for (int u = 0; u < RT; ++u) {
      const int r = u * TY + ty;
      if (r < rows) {
        long long row = (long long)(it.outAddr & kRowMask) + r;
        if (sp == kSpaceY && a.outPerm) row = a.outPerm[row];
        T *dst = reinterpret_cast<T *>(a.base[sp]) + row * ldOut + col0;
#pragma unroll
        for (int c = 0; c < CT; ++c) {
          if (colOk[c]) {
            T v = acc[u][c];
            if (it.flags & kItemAccumulate) v = S::add(v, dst[32 * c]);
            dst[32 * c] = v;
          }
        }
      }
    }
  }
}

// ---- many-RHS stage kernel, complex: DMMA (fp64 tensor-core) contraction ---------------------------
// With >= 2 RHS columns a stage is a batch of small dense contractions C(rows x nrhs) += A(rows x kc)
// X(kc x nrhs).  A CTA owns kWideQueuesPerCta queues x NC = 32 * NB RHS columns and is warp-specialised:
//
//   producer warp   walks the item / tile tables (the only global-memory reads of tables), allocates
//                   each tile a span of ONE byte ring in shared memory — packed A values followed by
//                   kc operand-row segments of NC columns (row stride padded by 32 bytes so that the
//                   B-fragment LDS.128 is conflict-free) — issues the TMA bulk copies for both, and
//                   publishes a 16-byte work entry {ring offset, shape, addresses, flags} under the
//                   same "full" mbarrier the copies complete on.  It runs up to kMmaBars tiles
//                   (~4 on average, several microseconds of DMMA work) ahead, which is what hides the
//                   DRAM latency of the operand rows: X does not fit in L2 at 256 RHS.
//   4 consumer warps  wait on "full", read the entry from shared memory, and own all <= 32 rows x the
//                   8 * NB columns [w * 8 NB, ...) as MB x NB m8n8k4 accumulator fragments, separately
//                   for the real and the imaginary part:
//                       Cr += Ar Xr + (-Ai) Xi        Ci += Ar Xi + Ai Xr     (4 DMMA / fragment / k4)
//                   then arrive on "empty" (which frees the ring span) and, after an item's last tile,
//                   store it.  No CTA-wide barrier and no global load sits between two tiles.
//
// Rows / k past the tile edge use clamped (valid) addresses; k past the edge multiplies by a zeroed A
// fragment; columns past nrhs are computed on stale shared memory and never stored.
__device__ __forceinline__ void dmma884(double (&c)[2], const double a, const double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
__device__ __forceinline__ void mbarArrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// One staged tile: C[mb][nb] += A(rows x kc) X(kc x 8 NB) for the first MBV row blocks.
template <int NB, int MBV, int MB>
__device__ __forceinline__ void mmaTile(double (&cr)[MB][NB][2], double (&ci)[MB][NB][2], const double2 *A,
                                        const unsigned char *Xs, const uint32_t XS, const int rows, const int kc,
                                        const int g, const int t4) {
  constexpr int ES = 16;
  const int ks = (kc + 3) >> 2;
  int ra[MBV];
#pragma unroll
  for (int mb = 0; mb < MBV; ++mb) ra[mb] = min(mb * 8 + g, rows - 1) * kc;
#pragma unroll 2
  for (int s = 0; s < ks; ++s) {
    const int k = s * 4 + t4;
    const bool kOk = k < kc;
    const int kcl = min(k, kc - 1);
    double2 xb[NB], av[MBV];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) xb[nb] = *reinterpret_cast<const double2 *>(Xs + (uint32_t)kcl * XS + nb * (8 * ES));
#pragma unroll
    for (int mb = 0; mb < MBV; ++mb) {
      av[mb] = A[ra[mb] + kcl];
      if (!kOk) av[mb] = make_double2(0.0, 0.0);
    }
#pragma unroll
    for (int mb = 0; mb < MBV; ++mb)
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) { dmma884(cr[mb][nb], av[mb].x, xb[nb].x); dmma884(ci[mb][nb], av[mb].x, xb[nb].y); }
#pragma unroll
    for (int mb = 0; mb < MBV; ++mb)
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) { dmma884(cr[mb][nb], -av[mb].y, xb[nb].y); dmma884(ci[mb][nb], av[mb].y, xb[nb].x); }
  }
}

// work entry flags (word 0, above the 20-bit ring offset)
constexpr uint32_t kEntFirst = 1u << 31, kEntLast = 1u << 30, kEntAccum = 1u << 29, kEntEnd = 1u << 28;
constexpr uint32_t kEntOffMask = (1u << 20) - 1;

template <int NB>
__global__ void __launch_bounds__(kMmaThreads) bfWideMmaStage(const __grid_constant__ StageLaunch a) {
  constexpr int ES = 16;
  constexpr int MB = kMaxRowsComplex / 8;
  constexpr int CW = kMmaThreads / 32 - 1; // consumer warps
  constexpr int NC = CW * NB * 8;
  constexpr uint32_t XS = NC * ES + 32;    // operand row stride in the ring
  constexpr uint32_t RING = kMmaRingBytes;
  // skip < size whenever the producer wraps, so 2 * size <= RING rules out a stall with nothing in flight
  static_assert(2 * (kTileBytes + kMaxKcComplex * XS) <= RING, "largest tile must fit the ring twice");
  static_assert(kMmaBars <= 32, "one lane per slot keeps its allocation size");

  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t *full = reinterpret_cast<uint64_t *>(smem + RING);
  uint64_t *empty = full + kMmaBars;
  uint4 *entries = reinterpret_cast<uint4 *>(empty + kMmaBars);
  const int tid = threadIdx.x;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < kMmaBars; ++s) { mbarInit(smemU32(&full[s]), 1); mbarInit(smemU32(&empty[s]), CW); }
    fenceBarrierInit();
  }
  __syncthreads();

  // grid = (column blocks, queue groups): the column blocks of one queue group are launched back to back, run
  // at the same time on different SMs and stream the SAME span of the pool at the same pace, so that span
  // comes out of DRAM once and out of L2 for the others (the pool is read with the default L2 policy when
  // there is more than one column block, evict-first when the CTA is the only reader)
  const int q0 = blockIdx.y * kWideQueuesPerCta;
  if (q0 >= a.numQueues) return;
  const int q1 = min(q0 + kWideQueuesPerCta, a.numQueues);
  const uint32_t iBeg = a.qItemBegin[q0], iEnd = a.qItemBegin[q1];
  if (iBeg == iEnd) return;

  const int lane = tid & 31, warp = tid >> 5;
  const long long c0 = (long long)a.colBegin + (long long)blockIdx.x * NC; // first column of the CTA

  if (warp == CW) {
    // ===== producer =====
    const uint64_t pol = policyEvictFirst();
    const uint32_t rowBytes = (uint32_t)min((long long)NC, (long long)a.nrhs - c0) * ES;
    const uint32_t ringBase = smemU32(smem);
    const unsigned char *pPool = a.pool + a.qPool16[q0] * 16ull;
    uint32_t pSeq = 0, oldest = 0, pOff = 0, inflight = 0;
    uint32_t allocReg = 0; // lane s: bytes to free when slot s drains
    // claim a slot and `size` contiguous ring bytes; returns the ring offset
    auto claim = [&](uint32_t size) -> uint32_t {
      const bool wrap = pOff + size > RING;
      const uint32_t skip = wrap ? RING - pOff : 0u;
      while (oldest + kMmaBars <= pSeq || inflight + skip + size > RING) {
        const uint32_t os = oldest % kMmaBars;
        mbarWait(smemU32(&empty[os]), (oldest / kMmaBars) & 1u);
        inflight -= __shfl_sync(0xffffffffu, allocReg, os);
        ++oldest;
      }
      const uint32_t start = wrap ? 0u : pOff;
      if (lane == (int)(pSeq % kMmaBars)) allocReg = skip + size;
      pOff = start + size;
      inflight += skip + size;
      return start;
    };
    auto publish = [&](uint4 e, uint32_t txBytes) {
      const uint32_t slot = pSeq % kMmaBars;
      if (lane == 0) {
        entries[slot] = e;
        mbarExpectTx(smemU32(&full[slot]), txBytes); // release: the entry is visible to whoever sees the phase
      }
      __syncwarp();
      ++pSeq;
    };
#pragma unroll 1
    for (uint32_t i = iBeg; i < iEnd; ++i) {
      const DevItem it = a.items[i];
      const uint32_t accum = (it.flags & kItemAccumulate) ? kEntAccum : 0u;
      if (it.numTiles == 0) {
        claim(0);
        publish(make_uint4(kEntFirst | kEntLast | accum, (uint32_t)it.rows << 16, 0u, it.outAddr), 0u);
        continue;
      }
#pragma unroll 1
      for (uint32_t t = it.tileBegin; t < it.tileBegin + it.numTiles; ++t) {
        const DevTile tl = a.tiles[t];
        const uint32_t bytes = tileBytes(ES, tl.rows, tl.kc);
        const uint32_t kEff = tl.kc >= kKcDiag ? tl.rows : tl.kc; // operand rows staged
        const uint32_t start = claim(bytes + kEff * XS);
        const uint32_t slot = pSeq % kMmaBars;
        const uint32_t bar = smemU32(&full[slot]);
        uint32_t w0 = start | accum;
        if (t == it.tileBegin) w0 |= kEntFirst;
        if (t + 1 == it.tileBegin + it.numTiles) w0 |= kEntLast;
        publish(make_uint4(w0, (uint32_t)tl.kc | ((uint32_t)tl.rows << 16), bytes, it.outAddr), bytes + kEff * rowBytes);
        if (lane == 0 && bytes) {
          if (gridDim.x == 1) bulkLoadHint(ringBase + start, pPool, bytes, bar, pol);
          else bulkLoad(ringBase + start, pPool, bytes, bar);
        }
        const uint32_t sp = tl.inAddr >> 30;
        const unsigned char *xsrc = a.base[sp] + ((long long)(tl.inAddr & kRowMask) * a.ld[sp] + c0) * ES;
        for (uint32_t r = lane; r < kEff; r += 32)
          bulkLoad(ringBase + start + bytes + r * XS, xsrc + (long long)r * a.ld[sp] * ES, rowBytes, bar);
        pPool += bytes;
      }
    }
    claim(0);
    publish(make_uint4(kEntEnd, 0u, 0u, 0u), 0u);
    return;
  }

  // ===== consumers =====
  const int g = lane >> 2, t4 = lane & 3;
  const long long cw = c0 + warp * (NB * 8); // first column of the warp
  const bool active = cw < a.nrhs;           // warp-uniform
  double cr[MB][NB][2], ci[MB][NB][2];
#pragma unroll 1
  for (uint32_t cSeq = 0;; ++cSeq) {
    const uint32_t slot = cSeq % kMmaBars;
    mbarWait(smemU32(&full[slot]), (cSeq / kMmaBars) & 1u);
    const uint4 e = entries[slot];
    if (e.x & kEntEnd) break;
    const int rows = (int)(e.y >> 16);
    const uint32_t kcRaw = e.y & 0xFFFFu;
    if (e.x & kEntFirst) {
#pragma unroll
      for (int mb = 0; mb < MB; ++mb)
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) { cr[mb][nb][0] = cr[mb][nb][1] = 0.0; ci[mb][nb][0] = ci[mb][nb][1] = 0.0; }
    }
    if (active && kcRaw != 0) { // kc == 0: an item no leaf covers (stores zeros)
      const unsigned char *T = smem + (e.x & kEntOffMask);
      const unsigned char *Xw = T + e.z + (uint32_t)(warp * (NB * 8)) * ES; // the warp's operand columns
      if (kcRaw >= kKcDiag) {
        // y[r] += d[r] * x[r]: operand row r sits at ring row r
        const double *dg = reinterpret_cast<const double *>(T);
#pragma unroll
        for (int mb = 0; mb < MB; ++mb) {
          const int r = mb * 8 + g;
          if (r < rows) {
            const double d = kcRaw == kKcDiag ? dg[r] : 1.0;
#pragma unroll
            for (int nb = 0; nb < NB; ++nb)
#pragma unroll
              for (int j = 0; j < 2; ++j) {
                const double2 xv = *reinterpret_cast<const double2 *>(Xw + (uint32_t)r * XS + (nb * 8 + 2 * t4 + j) * ES);
                cr[mb][nb][j] = ::fma(d, xv.x, cr[mb][nb][j]);
                ci[mb][nb][j] = ::fma(d, xv.y, ci[mb][nb][j]);
              }
          }
        }
      } else {
        const int kc = (int)kcRaw;
        const double2 *A = reinterpret_cast<const double2 *>(T);
        const unsigned char *Xs = Xw + g * ES;
        // exact fragment counts through a warp-uniform branch: a predicated-off DMMA still pays its
        // fixed issue delay, so rows <= 24 must not run the 4-block code
        switch ((rows + 7) >> 3) {
        case 1: mmaTile<NB, 1, MB>(cr, ci, A, Xs, XS, rows, kc, g, t4); break;
        case 2: mmaTile<NB, 2, MB>(cr, ci, A, Xs, XS, rows, kc, g, t4); break;
        case 3: mmaTile<NB, 3, MB>(cr, ci, A, Xs, XS, rows, kc, g, t4); break;
        default: mmaTile<NB, 4, MB>(cr, ci, A, Xs, XS, rows, kc, g, t4); break;
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbarArrive(smemU32(&empty[slot])); // the ring span and the entry may be reused

    if ((e.x & kEntLast) && active) {
      const uint32_t sp = e.w >> 30;
      const long long ldOut = a.ld[sp];
      const bool accum = (e.x & kEntAccum) != 0;
#pragma unroll
      for (int mb = 0; mb < MB; ++mb) {
        const int r = mb * 8 + g;
        if (r < rows) {
          long long row = (long long)(e.w & kRowMask) + r;
          if (sp == kSpaceY && a.outPerm) row = a.outPerm[row];
          double2 *dst = reinterpret_cast<double2 *>(a.base[sp]) + row * ldOut;
#pragma unroll
          for (int nb = 0; nb < NB; ++nb)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const long long c = cw + nb * 8 + 2 * t4 + j;
              if (c < a.nrhs) {
                double2 v = make_double2(cr[mb][nb][j], ci[mb][nb][j]);
                if (accum) { const double2 o = dst[c]; v.x += o.x; v.y += o.y; }
                dst[c] = v;
              }
            }
        }
      }
    }
  }
}

// ---- staging copy with the caller-side row permutation fused in ---------------------------------
__global__ void bfPermuteRows(const double *__restrict__ src, long long ldSrc, double *__restrict__ dst,
                              long long ldDst, const int32_t *__restrict__ perm, long long rows,
                              int rowDoubles, int elemDoubles) {
  // one thread per double of the row block; consecutive threads walk along a row
  const long long total = rows * rowDoubles;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long i = idx / rowDoubles, c = idx % rowDoubles;
    const long long o = perm ? (long long)perm[i] : i;
    dst[o * ldDst * elemDoubles + c] = src[i * ldSrc * elemDoubles + c];
  }
}

cudaError_t launchPermuteRows(const double *src, long long ldSrc, double *dst, long long ldDst,
                              const int32_t *perm, long long rows, int nrhs, int elemDoubles,
                              cudaStream_t stream) {
  const long long total = rows * nrhs * elemDoubles;
  if (total == 0) return cudaSuccess;
  const int threads = 256;
  const long long blocks = (total + threads - 1) / threads;
  bfPermuteRows<<<(unsigned)(blocks > 148 * 32 ? 148 * 32 : blocks), threads, 0, stream>>>(
      src, ldSrc, dst, ldDst, perm, rows, nrhs * elemDoubles, elemDoubles);
  return cudaGetLastError();
}

// ---- synthetic pool fill --------------------------------------------------------------------------
__global__ void bfFillSynthetic(int cplx, unsigned char *pool, const FillPiece *__restrict__ pieces,
                                long long numPieces, uint64_t seed) {
  for (long long pi = blockIdx.x; pi < numPieces; pi += gridDim.x) {
    const FillPiece p = pieces[pi];
    double *dst = reinterpret_cast<double *>(pool + p.pool16 * 16ull);
    if (p.kc == kKcIdentity) continue;
    if (p.kc == kKcDiag) {
      for (int r = threadIdx.x; r < p.rows; r += blockDim.x)
        dst[r] = 1.0 + 0.5 * syntheticValue(seed, p.rawIndex, p.row0 + r, 0, 1.0);
      continue;
    }
    const int n = (int)p.rows * (int)p.kc;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
      const int r = e / p.kc, c = e % p.kc;
      const long long pr = p.row0 + r, qc = p.col0 + c;
      const long long sr = p.swapRC ? qc : pr, sc = p.swapRC ? pr : qc;
      const long long lin = sr * p.storedCols + sc;
      if (cplx) {
        const double re = syntheticValue(seed, p.rawIndex, lin, 0, p.scale);
        const double im = syntheticValue(seed, p.rawIndex, lin, 1, p.scale);
        dst[2 * e] = re;
        dst[2 * e + 1] = p.conj ? -im : im;
      } else {
        dst[e] = syntheticValue(seed, p.rawIndex, lin, 0, p.scale);
      }
    }
  }
}

cudaError_t launchFillSynthetic(int dtype, unsigned char *pool, const FillPiece *pieces, long long numPieces,
                                uint64_t seed, cudaStream_t stream) {
  if (numPieces == 0) return cudaSuccess;
  const long long blocks = numPieces < 148 * 16 ? numPieces : 148 * 16;
  bfFillSynthetic<<<(unsigned)blocks, 128, 0, stream>>>(dtype == BF_CUDA_COMPLEX_F64, pool, pieces, numPieces, seed);
  return cudaGetLastError();
}

// ---- launchers ----------------------------------------------------------------------------------------
cudaError_t configureKernels() {
  cudaError_t e;
  e = cudaFuncSetAttribute(bfGemvStage<true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemvSmemBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(bfGemvStage<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemvSmemBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(bfGemvStage<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemvSmemBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(bfGemvStage<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemvSmemBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(bfGemvStage<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemvSmemBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(bfGemvStage<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemvSmemBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(bfWideMmaStage<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMmaSmemBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(bfWideMmaStage<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMmaSmemBytes);
  if (e != cudaSuccess) return e;
  return cudaSuccess;
}

cudaError_t launchGemvStage(int dtype, const StageLaunch &a, int /*numSMs*/, cudaStream_t stream) {
  const int ctas = (a.numQueues + kWarpsPerCta - 1) / kWarpsPerCta;
  if (ctas == 0) return cudaSuccess;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)ctas);
  cfg.blockDim = dim3(kWarpsPerCta * 32);
  cfg.dynamicSmemBytes = kGemvSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; // see griddepcontrol.wait in the kernel
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = a.pdl ? 1 : 0;
  return dtype == BF_CUDA_COMPLEX_F64 ? cudaLaunchKernelEx(&cfg, bfGemvStage<true, 0>, a) : cudaLaunchKernelEx(&cfg, bfGemvStage<false, 0>, a);
}

// One stage of one rank of a block-row sharded operator.  The "last warp posts the flags" count completes when all
// warps of the grid have finished; no warp waits on anything of its OWN grid or of a LATER stage (only on peers'
// previous stage), so CTAs need not be co-resident.
cudaError_t launchGemvShardStage(int dtype, const StageLaunch &a, cudaStream_t stream) {
  const int ctas = (a.numQueues + kWarpsPerCta - 1) / kWarpsPerCta;
  if (ctas == 0) return cudaSuccess;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)ctas);
  cfg.blockDim = dim3(kWarpsPerCta * 32);
  cfg.dynamicSmemBytes = kGemvSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = a.pdl ? 1 : 0;
  return dtype == BF_CUDA_COMPLEX_F64 ? cudaLaunchKernelEx(&cfg, bfGemvStage<true, 2>, a) : cudaLaunchKernelEx(&cfg, bfGemvStage<false, 2>, a);
}

// The whole 1-RHS apply as one cooperative launch (all CTAs co-resident: the readiness spins rely on it).
cudaError_t launchGemvFused(int dtype, const StageLaunch &a, cudaStream_t stream) {
  const int ctas = (a.numQueues + kWarpsPerCta - 1) / kWarpsPerCta;
  if (ctas == 0) return cudaSuccess;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)ctas);
  cfg.blockDim = dim3(kWarpsPerCta * 32);
  cfg.dynamicSmemBytes = kGemvSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return dtype == BF_CUDA_COMPLEX_F64 ? cudaLaunchKernelEx(&cfg, bfGemvStage<true, 1>, a) : cudaLaunchKernelEx(&cfg, bfGemvStage<false, 1>, a);
}

template <bool CPLX, int CT>
static cudaError_t launchWideT(const StageLaunch &a, int ncols, cudaStream_t stream) {
  const int ctas = (a.numQueues + kWideQueuesPerCta - 1) / kWideQueuesPerCta;
  dim3 grid((unsigned)ctas, (unsigned)((ncols + 32 * CT - 1) / (32 * CT)));
  bfWideStage<CPLX, CT><<<grid, kWideThreads, kWideSmemBytes, stream>>>(a);
  return cudaGetLastError();
}

template <int NB>
static cudaError_t launchWideMmaT(const StageLaunch &a, int ncols, cudaStream_t stream) {
  constexpr int NC = (kMmaThreads / 32 - 1) * NB * 8;
  const int ctas = (a.numQueues + kWideQueuesPerCta - 1) / kWideQueuesPerCta;
  dim3 grid((unsigned)((ncols + NC - 1) / NC), (unsigned)ctas);
  bfWideMmaStage<NB><<<grid, kMmaThreads, kMmaSmemBytes, stream>>>(a);
  return cudaGetLastError();
}

cudaError_t launchWideStage(int dtype, const StageLaunch &a, int /*numSMs*/, cudaStream_t stream) {
  const int ncols = a.nrhs - a.colBegin;
  if (ncols <= 0 || a.numQueues == 0) return cudaSuccess;
  if (dtype == BF_CUDA_COMPLEX_F64) {
    if (ncols > 32) return launchWideMmaT<2>(a, ncols, stream);
    return launchWideMmaT<1>(a, ncols, stream);
  }
  if (ncols > 64) return launchWideT<false, 4>(a, ncols, stream);
  if (ncols > 32) return launchWideT<false, 2>(a, ncols, stream);
  return launchWideT<false, 1>(a, ncols, stream);
}

} // namespace bfc
