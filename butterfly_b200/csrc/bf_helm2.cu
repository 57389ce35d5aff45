// bf_helm2.cu — the values of a fac_helm2 operator computed on the GPU, straight into the plan's packed
// pool.  Replaces, for a recorded operator (bf_helm2_record.c), the host work of
//   bfHelm2GetKernelMatrix       src/helm2.c:93-319    Z[i][j] = G(tgt_i, src_j), one Hankel function each
//   bfHelm2GetReexpansionMatrix  src/helm2.c:321-365   Z_equiv \ Z_orig
//   bfMatDenseComplexDenseComplexLstSq  src/mat_dense_complex.c:1768-1850   truncated-SVD least squares,
//                                       keep sigma_i >= max(m, n) eps sigma_0 + eps
//
// Two kernels:
//   bfHelm2Pinv   one CTA per DISTINCT Z_equiv (half of the re-expansion leaves share theirs with another
//                 leaf): builds Z_equiv in shared memory, orthogonalises its ROWS by one-sided Jacobi
//                 rotations (G Z_equiv = Sigma Q^H, G unitary, accumulated), applies the reference's
//                 truncation rule to the row norms sigma_i and stores the two factors of the pseudo-inverse,
//                 Q (kept columns) and Sigma^-1 G (kept rows).
//   bfHelm2Fill   one CTA per pool tile.  Kernel-matrix leaves: one Hankel function per element.
//                 Re-expansion leaves: X = Q (Sigma^-1 G Z_orig) for the tile's rows and columns, Z_orig
//                 evaluated on the fly and never stored.  The product is applied in factored form, small
//                 singular directions scaled BEFORE they are mixed (as the reference does with U_k^H, S_k^-1,
//                 V_k), so that an error of eps / sigma_i stays in direction i.
// Values land in the tile layout of every built direction (N / T / H): no staging copy, no host upload.
//
// Parity.  Kernel-matrix leaves agree with the reference to a few ulp of |H0| (CUDA's j0 / y0 against
// glibc's).  Re-expansion leaves are solutions of numerically rank-deficient least-squares problems
// (cond(Z_equiv) ~ 1e15): two LAPACK builds already differ from each other in the third digit of X there,
// while Z_equiv X - Z_orig agrees to 1e-14.  What is pinned (tests/test_gpu_helm2.py) is therefore what the
// reference itself checks (examples/simple/bf_all_blocks.c:149-153): the operator against the dense kernel
// matrix, next to the reference-built operator's own error on the same input.
#include <cuda_runtime.h>
#include <math_constants.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <tuple>
#include <vector>

#include "bf_internal.hpp"

using namespace bfc;

namespace {

constexpr int kFillThreads = 256;
constexpr int kTChunk = 32;      // targets per Hankel chunk
constexpr int kIBlock = 128;     // kept singular directions per pass (two threads each: half the columns)
constexpr int kMaxSweeps = 40;
constexpr int kMaxT = 4096;      // row norms and the compaction map live in shared memory

struct DevRecipe {
  int32_t kind, layerPot;
  double k;
  double2 alpha, beta;
  int64_t src, tgt, srcNormals, tgtNormals; // pair indices, -1 = no normals
  int32_t numSrc, numTgt;
  int32_t problem; // re-expansion: index of the pseudo-inverse
  int32_t pad;
};

struct DevProblem {
  int32_t T, ne;   // Z_equiv is T x ne (targets x equivalent sources)
  int32_t layerPot, threads;
  double k;
  double2 alpha, beta;
  int64_t tgt, equiv, equivNormals;
  int64_t qOff;    // Qn  [kk][ne]  : row i = kept row i of Sigma^-1 (G Z_equiv)  (= column i of Q, conjugated)
  int64_t gOff;    // GsT [T][T]    : GsT[t][i] = G[keep_i][t] / sigma_i
};

struct FillTile {
  uint64_t dst;       // device address of the tile (any direction's pool)
  int32_t recipe;
  int32_t row0, col0; // origin of the piece in the direction's acting matrix
  uint16_t rows, kc;
  uint8_t swapRC, conj;
  uint16_t pad;
};

__device__ __forceinline__ double2 cmul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ void cfma(double2 &acc, double2 a, double2 b) {
  acc.x = fma(a.x, b.x, acc.x); acc.x = fma(-a.y, b.y, acc.x);
  acc.y = fma(a.x, b.y, acc.y); acc.y = fma(a.y, b.x, acc.y);
}

// ---- Hankel functions (bf_hankel.inl) -----------------------------------------------------------------
#define BF_HANKEL_FN __device__
#define BF_HANKEL_TABLE __constant__
#include "bf_hankel.inl"

// One entry of a Helmholtz layer-potential kernel matrix (src/helm2.c:38-91 and :93-279): rows are targets,
// columns are sources; coincident points give 0 (src/helm2.c:113).  (i / 4) (J + i Y) = (-Y + i J) / 4.
__device__ double2 helmValue(int layerPot, double k, double2 alpha, double2 beta, double2 xs, double2 xt,
                             const double2 *ns, const double2 *nt) {
  const double r = hypot(xs.x - xt.x, xs.y - xt.y); // bfPoints2PairwiseDists, src/points.c:498
  if (r == 0) return make_double2(0.0, 0.0);
  const double kr = k * r;
  double2 S = make_double2(0.0, 0.0), D = make_double2(0.0, 0.0);
  if (layerPot == BF_CUDA_LAYER_POT_SINGLE || layerPot == BF_CUDA_LAYER_POT_COMBINED_FIELD) {
    const double2 h = hankel0(kr);
    S = make_double2(-0.25 * h.y, 0.25 * h.x);
  }
  if (layerPot != BF_CUDA_LAYER_POT_SINGLE) {
    // (i / 4) k H1(k r) / r * n . (x_tgt - x_src), n = source normal (D) or target normal (S')
    const double2 n = layerPot == BF_CUDA_LAYER_POT_PV_NORMAL_DERIV_SINGLE ? *nt : *ns;
    const double dot = n.x * (xt.x - xs.x) + n.y * (xt.y - xs.y);
    const double s = 0.25 * k / r * dot;
    const double2 h = hankel1(kr);
    D = make_double2(-s * h.y, s * h.x);
  }
  if (layerPot == BF_CUDA_LAYER_POT_SINGLE) return S;
  if (layerPot == BF_CUDA_LAYER_POT_COMBINED_FIELD) {
    const double2 a = cmul(alpha, S), b = cmul(beta, D);
    return make_double2(a.x + b.x, a.y + b.y);
  }
  return D;
}

// ---- pseudo-inverse factors of one Z_equiv ----------------------------------------------------------
// 64 registers per thread in every instantiation (the Jacobi loop needs ~40; only the one-off Hankel
// evaluation of Z_equiv spills): occupancy of the small classes is what the kernel's time depends on.
template <int THREADS>
__global__ void __launch_bounds__(THREADS, 1024 / THREADS)
bfHelm2Pinv(const DevProblem *__restrict__ problems, const int32_t *__restrict__ order, int numProblems,
            const double2 *__restrict__ pairs, double2 *__restrict__ factors, double2 *__restrict__ work,
            int32_t *__restrict__ kkOut, int32_t *__restrict__ sweepsOut, int smemA, int smemG,
            long long workStride, double cutScale, double tolScale, int usePrecond, double rankScale) {
  extern __shared__ __align__(16) unsigned char smemRaw[];
  __shared__ int sRot;
  __shared__ int sKept;
  for (int pi = blockIdx.x; pi < numProblems; pi += gridDim.x) {
    const int prob = order[pi];
    const DevProblem P = problems[prob];
    const int T = P.T, ne = P.ne;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nWarps = blockDim.x >> 5;
    // layout: [A: T x ne] [G: T x T] in shared memory when they fit, else in the global scratch
    unsigned char *sp = smemRaw;
    double2 *A, *G;
    double2 *gw = work + (size_t)blockIdx.x * (size_t)workStride; // per-CTA global scratch (unused when both fit)
    if (smemA) { A = reinterpret_cast<double2 *>(sp); sp += (size_t)T * ne * sizeof(double2); }
    else { A = gw; gw += (size_t)T * ne; }
    if (smemG) { G = reinterpret_cast<double2 *>(sp); sp += (size_t)T * T * sizeof(double2); }
    else G = gw;
    double2 *vbuf = reinterpret_cast<double2 *>(sp);        // Householder vector            [T]
    double *s2 = reinterpret_cast<double *>(vbuf + T);      // row norms^2 (singular values) [T]
    double *cn = s2 + T;                                    // trailing column norms^2       [ne]
    int *pos = reinterpret_cast<int *>(cn + ne);            // compaction map                [T]
    int *cperm = pos + T;                                   // column permutation of the QR  [ne]

    const double2 *tg = pairs + P.tgt, *eq = pairs + P.equiv;
    const double2 *en = P.equivNormals >= 0 ? pairs + P.equivNormals : nullptr;
    for (int e = tid; e < T * ne; e += blockDim.x) {
      const int t = e / ne, q = e - t * ne;
      A[e] = helmValue(P.layerPot, P.k, P.alpha, P.beta, eq[q], tg[t], en ? en + q : nullptr, nullptr);
    }
    for (int e = tid; e < T * T; e += blockDim.x) {
      const int t = e / T, u = e - t * T;
      G[e] = make_double2(t == u ? 1.0 : 0.0, 0.0);
    }
    for (int c = tid; c < ne; c += blockDim.x) cperm[c] = c;
    __syncthreads();

    // Preconditioner (Drmac & Veselic): Householder QR with column pivoting, Z_equiv P = Q_h R, G <- Q_h^H.
    // The rows of R are strongly graded, and Jacobi on them converges quadratically after ~4 sweeps instead
    // of ~16 (cond(Z_equiv) ~ 1e15 costs plain Jacobi about one sweep per decade).  Everything is "one thread
    // per column, loop over rows": contiguous in the row-major layout.
    // Rows of R beyond its numerical rank never reach the sweeps.  With column pivoting every column of the
    // trailing block is bounded by the pivot's norm |r_kk|, and for these matrices the pivots fall to a plateau
    // of pure rounding noise (~p eps |r_00|, the accumulated error of p reflections — the same size as the
    // truncation threshold max(m, n) eps sigma_0) after ~0.4 T steps and stay there.  The factorization stops
    // when the pivot is below rankScale (4) times that threshold: what is left is noise whose singular
    // directions the reference's rule cuts or is indifferent to.  Measured: the operator's error
    // against the dense kernel is unchanged in the fourth digit for rankScale 0 / 0.25 / 1 / 2 / 4, while the
    // largest launch class (scratch in L2 / HBM) goes from 11.1 s to 7.0 s at N = 524 288.
    int Teff = T;
    if (usePrecond) {
      const int steps = T - 1 < ne ? T - 1 : ne;
      Teff = T < ne ? T : ne; // rows below the n_e-th are exactly zero after a full factorization
      double first2 = 0.0;
      for (int k = 0; k < steps; ++k) {
        for (int c = k + tid; c < ne; c += blockDim.x) {
          double a = 0;
          for (int r = k; r < T; ++r) { const double2 x = A[(size_t)r * ne + c]; a = fma(x.x, x.x, fma(x.y, x.y, a)); }
          cn[c] = a;
        }
        __syncthreads();
        if (warp == 0) { // pivot = trailing column of largest norm (first one on ties)
          double best = -1.0;
          int bi = k;
          for (int c = k + lane; c < ne; c += 32) if (cn[c] > best) { best = cn[c]; bi = c; }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
          }
          if (lane == 0) { sRot = bi; s2[0] = best; }
        }
        __syncthreads();
        const int piv = sRot;
        const double nrm2 = s2[0];
        if (k == 0) first2 = nrm2;
        // the trailing block is negligible (or zero, or so small that 1 / ||x||^2 would overflow)
        const double small = rankScale * (double)(T > ne ? T : ne) * 2.220446049250313e-16;
        if (!(nrm2 > 1e-280) || nrm2 < small * small * first2) { Teff = k; break; }
        if (piv != k) {
          for (int r = tid; r < T; r += blockDim.x) {
            const double2 t0 = A[(size_t)r * ne + k];
            A[(size_t)r * ne + k] = A[(size_t)r * ne + piv];
            A[(size_t)r * ne + piv] = t0;
          }
          if (tid == 0) { const int t0 = cperm[k]; cperm[k] = cperm[piv]; cperm[piv] = t0; }
        }
        __syncthreads();
        // H = I - tau v v^H with v = x - alpha e_1, alpha = -e^{i arg x_0} ||x||:  H x = alpha e_1
        const double2 x0 = A[(size_t)k * ne + k];
        const double nrm = sqrt(nrm2), ax0 = sqrt(x0.x * x0.x + x0.y * x0.y);
        const double2 ph = ax0 > 0 ? make_double2(x0.x / ax0, x0.y / ax0) : make_double2(1.0, 0.0);
        const double tau = 1.0 / (nrm * (nrm + ax0));
        for (int r = k + tid; r < T; r += blockDim.x) {
          double2 v = A[(size_t)r * ne + k];
          if (r == k) v = make_double2(ph.x * (ax0 + nrm), ph.y * (ax0 + nrm));
          vbuf[r] = v;
        }
        __syncthreads();
        // columns k+1.. of A and all columns of G:  M[k:, c] -= tau v (v^H M[k:, c])
        for (int c = tid; c < (ne - k - 1) + T; c += blockDim.x) {
          const bool inA = c < ne - k - 1;
          double2 *M = inA ? A + (k + 1 + c) : G + (c - (ne - k - 1));
          const size_t ld = inA ? (size_t)ne : (size_t)T;
          double2 w = make_double2(0.0, 0.0);
          for (int r = k; r < T; ++r) { const double2 v = vbuf[r]; cfma(w, make_double2(v.x, -v.y), M[(size_t)r * ld]); }
          w.x *= tau; w.y *= tau;
          for (int r = k; r < T; ++r) {
            double2 y = M[(size_t)r * ld];
            cfma(y, make_double2(-vbuf[r].x, -vbuf[r].y), w);
            M[(size_t)r * ld] = y;
          }
        }
        // column k itself becomes (alpha, 0, ..., 0)
        for (int r = k + tid; r < T; r += blockDim.x)
          A[(size_t)r * ne + k] = r == k ? make_double2(-ph.x * nrm, -ph.y * nrm) : make_double2(0.0, 0.0);
        __syncthreads();
      }
      __syncthreads();
    }

    // one-sided (Hestenes) Jacobi on the rows, round-robin tournament order: every round holds disjoint pairs
    const int m = (Teff + 1) & ~1, half = m >> 1;
    const double tol = tolScale * sqrt((double)ne) * 2.220446049250313e-16;
    int sweep = 0;
    for (; sweep < kMaxSweeps; ++sweep) {
      if (tid == 0) sRot = 0;
      __syncthreads();
      for (int rd = 0; rd < m - 1; ++rd) {
        for (int q = warp; q < half; q += nWarps) {
          int i, j;
          if (q == 0) { i = m - 1; j = rd; }
          else { i = (rd + q) % (m - 1); j = (rd - q + (m - 1)) % (m - 1); }
          if (i >= Teff || j >= Teff) continue; // the bye of an odd tournament
          if (i > j) { const int s = i; i = j; j = s; }
          double2 *ai = A + (size_t)i * ne, *aj = A + (size_t)j * ne;
          double a = 0, b = 0, cr = 0, ci = 0; // ||a_i||^2, ||a_j||^2, sum a_i conj(a_j)
          for (int c = lane; c < ne; c += 32) {
            const double2 x = ai[c], y = aj[c];
            a = fma(x.x, x.x, fma(x.y, x.y, a));
            b = fma(y.x, y.x, fma(y.y, y.y, b));
            cr = fma(x.x, y.x, fma(x.y, y.y, cr));
            ci = fma(x.y, y.x, fma(-x.x, y.y, ci));
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
            cr += __shfl_xor_sync(0xffffffffu, cr, o);
            ci += __shfl_xor_sync(0xffffffffu, ci, o);
          }
          const double ac2 = cr * cr + ci * ci;
          if (!(ac2 > tol * tol * a * b)) continue; // converged pair (also skips zero rows): no sqrt, no divide
          if (lane == 0) atomicAdd(&sRot, 1);
          // rotate [a_i; e^{i theta} a_j] by the real Jacobi angle that zeroes their inner product
          const double iac = rsqrt(ac2);
          const double2 ph = make_double2(cr * iac, ci * iac);
          const double zeta = 0.5 * (b - a) * iac;
          const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
          const double cs = rsqrt(1.0 + t * t), sn = cs * t;
          for (int c = lane; c < ne; c += 32) {
            const double2 x = ai[c], y = cmul(ph, aj[c]);
            ai[c] = make_double2(cs * x.x - sn * y.x, cs * x.y - sn * y.y);
            aj[c] = make_double2(sn * x.x + cs * y.x, sn * x.y + cs * y.y);
          }
          double2 *gi = G + (size_t)i * T, *gj = G + (size_t)j * T;
          for (int c = lane; c < T; c += 32) {
            const double2 x = gi[c], y = cmul(ph, gj[c]);
            gi[c] = make_double2(cs * x.x - sn * y.x, cs * x.y - sn * y.y);
            gj[c] = make_double2(sn * x.x + cs * y.x, sn * x.y + cs * y.y);
          }
        }
        __syncthreads();
      }
      const int rot = sRot;
      __syncthreads();
      if (rot == 0) { ++sweep; break; }
    }

    // singular values = row norms; the reference's truncation (src/mat_dense_complex.c:1796-1806)
    for (int i = warp; i < Teff; i += nWarps) {
      double a = 0;
      for (int c = lane; c < ne; c += 32) { const double2 x = A[(size_t)i * ne + c]; a = fma(x.x, x.x, fma(x.y, x.y, a)); }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
      if (lane == 0) s2[i] = a;
    }
    __syncthreads();
    if (tid == 0) {
      double mx = 0;
      for (int i = 0; i < Teff; ++i) mx = fmax(mx, s2[i]);
      const double eps = 2.220446049250313e-16;
      const double cut = cutScale * ((double)(T > ne ? T : ne) * eps * sqrt(mx) + eps);
      int kk = 0;
      const int cap = T < ne ? T : ne; // rank <= min(m, n): the reference only has that many sigma
      // keep the largest first when more than `cap` rows pass (cannot happen once the sweeps converged)
      for (int i = 0; i < T; ++i) {
        const bool keep = i < Teff && sqrt(s2[i]) >= cut && kk < cap;
        pos[i] = keep ? kk++ : -1;
      }
      sKept = kk;
      kkOut[prob] = kk;
      sweepsOut[prob] = sweep;
    }
    __syncthreads();
    const int kk = sKept;
    double2 *Qn = factors + P.qOff, *GsT = factors + P.gOff;
    for (int e = tid; e < T * ne; e += blockDim.x) {
      const int i = e / ne, c = e - i * ne;
      if (pos[i] < 0) continue;
      const double inv = 1.0 / sqrt(s2[i]);
      const double2 x = A[e];
      Qn[(size_t)pos[i] * ne + cperm[c]] = make_double2(x.x * inv, x.y * inv); // undo the QR's column pivoting
    }
    for (int e = tid; e < T * T; e += blockDim.x) {
      const int t = e / T, i = e - t * T; // write GsT[t][pos[i]]: coalesced along i
      if (pos[i] < 0) continue;
      const double inv = 1.0 / sqrt(s2[i]);
      const double2 x = G[(size_t)i * T + t];
      GsT[(size_t)t * T + pos[i]] = make_double2(x.x * inv, x.y * inv);
    }
    (void)kk;
    __syncthreads();
  }
}

// ---- pool fill -----------------------------------------------------------------------------------------
template <int NC>
__device__ __forceinline__ void reexpTile(const FillTile &ft, const DevRecipe &rc, const DevProblem &P, int kk,
                                          const double2 *__restrict__ pairs, const double2 *__restrict__ factors,
                                          double2 *dst, double2 *Ks, double2 *Cs) {
  const int tid = threadIdx.x;
  const int T = P.T, ne = P.ne;
  // stored sub-block of X = Z_equiv \ Z_orig this tile holds: rows [R0, R0 + nr), columns [C0, C0 + nc)
  const int nr = ft.swapRC ? ft.kc : ft.rows, nc = ft.swapRC ? ft.rows : ft.kc;
  const int R0 = ft.swapRC ? ft.col0 : ft.row0, C0 = ft.swapRC ? ft.row0 : ft.col0;
  const double2 *tg = pairs + rc.tgt, *sr = pairs + rc.src + C0;
  const double2 *sn = rc.srcNormals >= 0 ? pairs + rc.srcNormals + C0 : nullptr;
  const double2 *Qn = factors + P.qOff, *GsT = factors + P.gOff;
  const int n = (int)ft.rows * (int)ft.kc;
  constexpr int kOut = 512 / kFillThreads; // a tile holds at most 512 entries
  constexpr int HC = NC / 2;               // columns per thread in the first product
  const int il = tid >> 1, c0 = (tid & 1) * HC;
  double2 out[kOut];
#pragma unroll
  for (int j = 0; j < kOut; ++j) out[j] = make_double2(0.0, 0.0);
  for (int i0 = 0; i0 < kk; i0 += kIBlock) {
    const int i = i0 + il;
    double2 acc[HC];
#pragma unroll
    for (int c = 0; c < HC; ++c) acc[c] = make_double2(0.0, 0.0);
    for (int t0 = 0; t0 < T; t0 += kTChunk) {
      __syncthreads();
      for (int e = tid; e < kTChunk * NC; e += kFillThreads) {
        const int tt = e / NC, c = e - tt * NC, t = t0 + tt;
        double2 v = make_double2(0.0, 0.0);
        if (t < T && c < nc) v = helmValue(rc.layerPot, rc.k, rc.alpha, rc.beta, sr[c], tg[t], sn ? sn + c : nullptr, nullptr);
        Ks[e] = v;
      }
      __syncthreads();
      if (i < kk) {
        const int tEnd = min(kTChunk, T - t0);
        for (int tt = 0; tt < tEnd; ++tt) {
          const double2 g = GsT[(size_t)(t0 + tt) * T + i];
#pragma unroll
          for (int c = 0; c < HC; ++c) cfma(acc[c], g, Ks[tt * NC + c0 + c]);
        }
      }
    }
    __syncthreads();
    // row stride NC + 2 and the second half shifted by one element: the 8 lanes of a 128-bit store phase (4 rows x
    // 2 halves) land in 8 different 16-byte bank groups (a stride of NC * 16 B put all of them in one)
#pragma unroll
    for (int c = 0; c < HC; ++c) Cs[il * (NC + 2) + (tid & 1) * (HC + 1) + c] = i < kk ? acc[c] : make_double2(0.0, 0.0);
    __syncthreads();
    const int ib = min(kIBlock, kk - i0);
#pragma unroll
    for (int j = 0; j < kOut; ++j) {
      const int e = tid + j * kFillThreads;
      if (e >= n) continue;
      const int ra = e / ft.kc, ca = e - ra * ft.kc;
      const int rs = ft.swapRC ? ca : ra, cs = ft.swapRC ? ra : ca; // (row, column) inside the stored sub-block
      double2 s = out[j];
      for (int ii = 0; ii < ib; ++ii) {
        const double2 q = Qn[(size_t)(i0 + ii) * ne + R0 + rs];
        cfma(s, make_double2(q.x, -q.y), Cs[ii * (NC + 2) + cs + (cs >= HC ? 1 : 0)]);
      }
      out[j] = s;
    }
  }
  (void)nr;
#pragma unroll
  for (int j = 0; j < kOut; ++j) {
    const int e = tid + j * kFillThreads;
    if (e < n) dst[e] = make_double2(out[j].x, ft.conj ? -out[j].y : out[j].y);
  }
}

__global__ void __launch_bounds__(kFillThreads, 2)
bfHelm2Fill(const FillTile *__restrict__ tiles, long long numTiles,
            const DevRecipe *__restrict__ recipes, const DevProblem *__restrict__ problems,
            const int32_t *__restrict__ kkOf, const double2 *__restrict__ pairs,
            const double2 *__restrict__ factors, unsigned long long *counter) {
  extern __shared__ __align__(16) unsigned char smemRaw[];
  double2 *Ks = reinterpret_cast<double2 *>(smemRaw);            // [kTChunk][32]
  double2 *Cs = Ks + kTChunk * 32;                               // [kIBlock][32 + 2]
  __shared__ long long sTile;
  const int tid = threadIdx.x;
  for (;;) {
    __syncthreads();
    if (tid == 0) sTile = (long long)atomicAdd(counter, 1ull);
    __syncthreads();
    const long long ti = sTile;
    if (ti >= numTiles) break;
    const FillTile ft = tiles[ti];
    const DevRecipe rc = recipes[ft.recipe];
    double2 *dst = reinterpret_cast<double2 *>(ft.dst);
    const int n = (int)ft.rows * (int)ft.kc;
    if (rc.kind == BF_CUDA_HELM2_KERNEL) {
      const double2 *tg = pairs + rc.tgt, *sr = pairs + rc.src;
      const double2 *sn = rc.srcNormals >= 0 ? pairs + rc.srcNormals : nullptr;
      const double2 *tn = rc.tgtNormals >= 0 ? pairs + rc.tgtNormals : nullptr;
      for (int e = tid; e < n; e += kFillThreads) {
        const int ra = e / ft.kc, ca = e - ra * ft.kc;
        const int p = ft.row0 + ra, q = ft.col0 + ca;
        const int rs = ft.swapRC ? q : p, cs = ft.swapRC ? p : q; // stored (target, source)
        const double2 v = helmValue(rc.layerPot, rc.k, rc.alpha, rc.beta, sr[cs], tg[rs], sn ? sn + cs : nullptr,
                                    tn ? tn + rs : nullptr);
        dst[e] = make_double2(v.x, ft.conj ? -v.y : v.y);
      }
      continue;
    }
    const DevProblem P = problems[rc.problem];
    const int kk = kkOf[rc.problem];
    const int nc = ft.swapRC ? ft.rows : ft.kc;
    if (nc <= 8) reexpTile<8>(ft, rc, P, kk, pairs, factors, dst, Ks, Cs);
    else if (nc <= 16) reexpTile<16>(ft, rc, P, kk, pairs, factors, dst, Ks, Cs);
    else reexpTile<32>(ft, rc, P, kk, pairs, factors, dst, Ks, Cs);
  }
}

constexpr int kFillSmemBytes = (kTChunk * 32 + kIBlock * (32 + 2)) * (int)sizeof(double2);

// ---- host side ---------------------------------------------------------------------------------------
struct DeviceBuffers {
  std::vector<void *> ptrs;
  ~DeviceBuffers() { for (void *p : ptrs) cudaFree(p); }
  template <typename T> cudaError_t alloc(T **p, size_t count) {
    cudaError_t e = cudaMalloc(reinterpret_cast<void **>(p), std::max<size_t>(count, 1) * sizeof(T));
    if (e == cudaSuccess) ptrs.push_back(*p);
    return e;
  }
};

struct Prepared {
  std::vector<DevRecipe> recipes;
  std::vector<DevProblem> problems;
  size_t factorElems = 0;
  // launch classes of the pseudo-inverse kernel; workStride = global scratch elements one CTA needs
  struct Class { int smemA, smemG, threads; size_t smemBytes; std::vector<int32_t> order; size_t workStride; };
  std::vector<Class> classes;
};

BfCudaStatus checkRange(int64_t first, int64_t count, int64_t numPairs, const char *what) {
  if (first < 0 || count < 0 || first + count > numPairs) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, std::string("recipe ") + what + " outside the pair pool");
  return BF_CUDA_OK;
}

BfCudaStatus prepare(const BfCudaHelm2Build &b, const std::vector<char> &used, Prepared &out, size_t smemLimit) {
  out.recipes.resize((size_t)b.numRecipes);
  using Key = std::tuple<int64_t, int64_t, int64_t, int64_t, int64_t, int32_t, uint64_t, uint64_t, uint64_t, uint64_t, uint64_t>;
  std::map<Key, int32_t> seen;
  auto bits = [](double d) { uint64_t u; std::memcpy(&u, &d, 8); return u; };
  for (int64_t r = 0; r < b.numRecipes; ++r) {
    const BfCudaHelm2Recipe &c = b.recipes[r];
    DevRecipe &d = out.recipes[(size_t)r];
    std::memset(&d, 0, sizeof d);
    d.kind = c.kind; d.layerPot = c.layerPot; d.k = c.k;
    d.alpha = make_double2(c.alpha[0], c.alpha[1]); d.beta = make_double2(c.beta[0], c.beta[1]);
    d.src = c.src; d.tgt = c.tgt; d.srcNormals = c.srcNormals; d.tgtNormals = c.tgtNormals;
    d.numSrc = (int32_t)c.numSrc; d.numTgt = (int32_t)c.numTgt; d.problem = -1;
    if (!used[(size_t)r]) continue;
    if (c.kind != BF_CUDA_HELM2_KERNEL && c.kind != BF_CUDA_HELM2_REEXPANSION) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "recipe kind");
    const bool needSrcN = c.layerPot == BF_CUDA_LAYER_POT_PV_DOUBLE || c.layerPot == BF_CUDA_LAYER_POT_COMBINED_FIELD;
    const bool needTgtN = c.layerPot == BF_CUDA_LAYER_POT_PV_NORMAL_DERIV_SINGLE;
    if (c.layerPot != BF_CUDA_LAYER_POT_SINGLE && !needSrcN && !needTgtN) return fail(BF_CUDA_ERROR_NOT_IMPLEMENTED, "layer potential (src/helm2.c:303-315 implements S, D, S' and the combined field)");
    if (!(c.k > 0)) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "wavenumber must be positive (src/helm2.c:138)");
    BfCudaStatus st = checkRange(c.src, c.numSrc, b.numPairs, "sources");
    if (st == BF_CUDA_OK) st = checkRange(c.tgt, c.numTgt, b.numPairs, "targets");
    if (st == BF_CUDA_OK && needSrcN) st = checkRange(c.srcNormals, c.numSrc, b.numPairs, "source normals");
    if (st != BF_CUDA_OK) return st;
    if (c.kind == BF_CUDA_HELM2_KERNEL) {
      if (needTgtN && (st = checkRange(c.tgtNormals, c.numTgt, b.numPairs, "target normals")) != BF_CUDA_OK) return st;
      if (c.rows != c.numTgt || c.cols != c.numSrc) return fail(BF_CUDA_ERROR_INCOMPATIBLE_SHAPES, "kernel recipe shape");
      continue;
    }
    if (needTgtN) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "no re-expansion for a potential that uses target normals (src/helm2.c:339-340)");
    if ((st = checkRange(c.equiv, c.numEquiv, b.numPairs, "equivalent sources")) != BF_CUDA_OK) return st;
    if (needSrcN && (st = checkRange(c.equivNormals, c.numEquiv, b.numPairs, "equivalent-source normals")) != BF_CUDA_OK) return st;
    if (c.rows != c.numEquiv || c.cols != c.numSrc) return fail(BF_CUDA_ERROR_INCOMPATIBLE_SHAPES, "re-expansion recipe shape");
    if (c.numTgt > kMaxT || c.numEquiv > kMaxT) return fail(BF_CUDA_ERROR_NOT_IMPLEMENTED, "re-expansion with more than 4096 proxy points");
    const Key key{c.equiv, c.numEquiv, c.tgt, c.numTgt, needSrcN ? c.equivNormals : -1, c.layerPot, bits(c.k),
                  bits(c.alpha[0]), bits(c.alpha[1]), bits(c.beta[0]), bits(c.beta[1])};
    auto it = seen.find(key);
    if (it == seen.end()) {
      DevProblem p;
      std::memset(&p, 0, sizeof p);
      p.T = (int32_t)c.numTgt; p.ne = (int32_t)c.numEquiv; p.layerPot = c.layerPot; p.k = c.k;
      p.alpha = d.alpha; p.beta = d.beta;
      p.tgt = c.tgt; p.equiv = c.equiv; p.equivNormals = needSrcN ? c.equivNormals : -1;
      const size_t T = (size_t)p.T, ne = (size_t)p.ne;
      p.qOff = (int64_t)out.factorElems; out.factorElems += T * ne;
      p.gOff = (int64_t)out.factorElems; out.factorElems += T * T;
      it = seen.emplace(key, (int32_t)out.problems.size()).first;
      out.problems.push_back(p);
    }
    d.problem = it->second;
  }
  // launch classes: (A and G in shared memory) by size, then (A only), then (neither)
  const size_t caps[] = {32 * 1024, 72 * 1024, 110 * 1024, smemLimit};
  out.classes.clear();
  const int classThreads[] = {256, 512, 1024, 1024}; // enough warps to hold a round's pairs (T / 2) at once
  for (size_t c = 0; c < 4; ++c) out.classes.push_back({1, 1, classThreads[c], caps[c], {}, 0});
  out.classes.push_back({1, 0, 1024, smemLimit, {}, 0});
  out.classes.push_back({0, 0, 1024, 0, {}, 0});
  for (size_t i = 0; i < out.problems.size(); ++i) {
    DevProblem &p = out.problems[i];
    const size_t T = (size_t)p.T, ne = (size_t)p.ne;
    const size_t aux = T * 28 + ne * 12 + 128, aBytes = T * ne * 16, gBytes = T * T * 16; // vbuf, s2, pos | cn, cperm
    size_t cls = 5;
    for (size_t c = 0; c < 4 && cls == 5; ++c) if (aBytes + gBytes + aux <= caps[c]) cls = c;
    if (cls == 5 && aBytes + aux <= smemLimit) cls = 4;
    if (cls == 4) out.classes[cls].workStride = std::max(out.classes[cls].workStride, T * T);
    if (cls == 5) out.classes[cls].workStride = std::max(out.classes[cls].workStride, T * ne + T * T);
    out.classes[cls].order.push_back((int32_t)i);
    if (cls == 5) out.classes[cls].smemBytes = std::max(out.classes[cls].smemBytes, aux);
  }
  return BF_CUDA_OK;
}

// Experiment knobs (default 1): scale of the truncation threshold / of the Jacobi convergence tolerance.
double envScale(const char *name) {
  const char *v = std::getenv(name);
  const double d = v ? std::atof(v) : 1.0;
  return d > 0 ? d : 1.0;
}

struct Uploaded {
  double2 *pairs = nullptr, *factors = nullptr;
  DevRecipe *recipes = nullptr;
  DevProblem *problems = nullptr;
  int32_t *kk = nullptr, *sweeps = nullptr;
  unsigned long long *counter = nullptr;
};

BfCudaStatus runPinv(const BfCudaHelm2Build &b, const Prepared &prep, DeviceBuffers &mem, Uploaded &u, int numSMs,
                     cudaStream_t stream, BfCudaHelm2Stats *stats) {
  BF_CUDA_TRY(mem.alloc(&u.pairs, (size_t)b.numPairs));
  BF_CUDA_TRY(cudaMemcpyAsync(u.pairs, b.pairs, (size_t)b.numPairs * sizeof(double2), cudaMemcpyHostToDevice, stream));
  BF_CUDA_TRY(mem.alloc(&u.recipes, prep.recipes.size()));
  BF_CUDA_TRY(cudaMemcpyAsync(u.recipes, prep.recipes.data(), prep.recipes.size() * sizeof(DevRecipe), cudaMemcpyHostToDevice, stream));
  const size_t nP = prep.problems.size();
  BF_CUDA_TRY(mem.alloc(&u.problems, nP));
  BF_CUDA_TRY(cudaMemcpyAsync(u.problems, prep.problems.data(), nP * sizeof(DevProblem), cudaMemcpyHostToDevice, stream));
  BF_CUDA_TRY(mem.alloc(&u.factors, prep.factorElems));
  BF_CUDA_TRY(mem.alloc(&u.kk, nP));
  BF_CUDA_TRY(mem.alloc(&u.sweeps, nP));
  BF_CUDA_TRY(mem.alloc(&u.counter, 1));
  BF_CUDA_TRY(cudaMemsetAsync(u.kk, 0, std::max<size_t>(nP, 1) * 4, stream));
  BF_CUDA_TRY(cudaMemsetAsync(u.sweeps, 0, std::max<size_t>(nP, 1) * 4, stream));
  cudaEvent_t e0, e1;
  BF_CUDA_TRY(cudaEventCreate(&e0));
  BF_CUDA_TRY(cudaEventCreate(&e1));
  BF_CUDA_TRY(cudaEventRecord(e0, stream));
  std::vector<cudaEvent_t> classEv(prep.classes.size() + 1, nullptr);
  for (auto &e : classEv) BF_CUDA_TRY(cudaEventCreate(&e));
  BF_CUDA_TRY(cudaEventRecord(classEv[0], stream));
  for (size_t ci = 0; ci < prep.classes.size(); ++ci) {
    const Prepared::Class &c = prep.classes[ci];
    if (c.order.empty()) { BF_CUDA_TRY(cudaEventRecord(classEv[ci + 1], stream)); continue; }
    int32_t *dOrder = nullptr;
    BF_CUDA_TRY(mem.alloc(&dOrder, c.order.size()));
    BF_CUDA_TRY(cudaMemcpyAsync(dOrder, c.order.data(), c.order.size() * 4, cudaMemcpyHostToDevice, stream));
    auto kernel = c.threads == 256 ? bfHelm2Pinv<256> : c.threads == 512 ? bfHelm2Pinv<512> : bfHelm2Pinv<1024>;
    BF_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smemBytes));
    // without this the driver may pick a small shared-memory carve-out and run ONE 32 KiB CTA per SM
    // (the class that keeps everything in global scratch wants the L1 instead)
    BF_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     c.smemA ? cudaSharedmemCarveoutMaxShared : cudaSharedmemCarveoutMaxL1));
    int perSM = 1;
    BF_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, kernel, c.threads, c.smemBytes));
    const size_t grid = std::min<size_t>(c.order.size(), (size_t)numSMs * (size_t)std::max(perSM, 1));
    double2 *work = nullptr;
    if (c.workStride) BF_CUDA_TRY(mem.alloc(&work, grid * c.workStride));
    kernel<<<(unsigned)grid, c.threads, c.smemBytes, stream>>>(u.problems, dOrder, (int)c.order.size(), u.pairs, u.factors,
                                                                    work, u.kk, u.sweeps, c.smemA, c.smemG, (long long)c.workStride,
                                                                    b.truncationScale > 0 ? b.truncationScale : envScale("BF_CUDA_HELM2_CUT_SCALE"),
                                                                    envScale("BF_CUDA_HELM2_TOL_SCALE"),
                                                                    std::getenv("BF_CUDA_HELM2_NO_QR") ? 0 : 1,
                                                                    std::getenv("BF_CUDA_HELM2_RANK_SCALE") ? std::atof(std::getenv("BF_CUDA_HELM2_RANK_SCALE")) : 4.0);
    BF_CUDA_TRY(cudaGetLastError());
    if (stats) { if (c.smemA && c.smemG) stats->reexpSmall += (int64_t)c.order.size(); else stats->reexpLarge += (int64_t)c.order.size(); }
    BF_CUDA_TRY(cudaEventRecord(classEv[ci + 1], stream));
  }
  BF_CUDA_TRY(cudaEventRecord(e1, stream));
  BF_CUDA_TRY(cudaEventSynchronize(e1));
  float ms = 0;
  BF_CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  for (size_t ci = 0; ci < prep.classes.size() && ci < 6; ++ci) {
    float cm = 0;
    if (stats && cudaEventElapsedTime(&cm, classEv[ci], classEv[ci + 1]) == cudaSuccess) {
      stats->pinvClassMs[ci] = cm;
      stats->pinvClassCount[ci] = (int64_t)prep.classes[ci].order.size();
    }
  }
  for (auto &e : classEv) cudaEventDestroy(e);
  if (stats) {
    stats->reexpMs = ms;
    std::vector<int32_t> sw(nP);
    if (nP) BF_CUDA_TRY(cudaMemcpy(sw.data(), u.sweeps, nP * 4, cudaMemcpyDeviceToHost));
    for (int32_t s : sw) {
      stats->maxSweeps = std::max<int64_t>(stats->maxSweeps, s);
      if (s >= kMaxSweeps) ++stats->notConverged;
    }
  }
  return BF_CUDA_OK;
}

BfCudaStatus runFill(const std::vector<FillTile> &tiles, const Uploaded &u, DeviceBuffers &mem, int numSMs,
                     cudaStream_t stream, float *ms) {
  if (tiles.empty()) return BF_CUDA_OK;
  FillTile *dTiles = nullptr;
  BF_CUDA_TRY(cudaMalloc(&dTiles, tiles.size() * sizeof(FillTile)));
  cudaError_t e = cudaMemcpyAsync(dTiles, tiles.data(), tiles.size() * sizeof(FillTile), cudaMemcpyHostToDevice, stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(u.counter, 0, sizeof(unsigned long long), stream);
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (e == cudaSuccess) e = cudaEventCreate(&e0);
  if (e == cudaSuccess) e = cudaEventCreate(&e1);
  if (e == cudaSuccess) {
    const size_t grid = std::min<size_t>(tiles.size(), (size_t)numSMs * 2);
    cudaEventRecord(e0, stream);
    bfHelm2Fill<<<(unsigned)grid, kFillThreads, kFillSmemBytes, stream>>>(dTiles, (long long)tiles.size(), u.recipes,
                                                                         u.problems, u.kk, u.pairs, u.factors, u.counter);
    e = cudaGetLastError();
    cudaEventRecord(e1, stream);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  if (e == cudaSuccess && ms) cudaEventElapsedTime(ms, e0, e1);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  cudaFree(dTiles);
  (void)mem;
  if (e != cudaSuccess) return fail(BF_CUDA_ERROR_RUNTIME, std::string("helm2 fill: ") + cudaGetErrorString(e));
  return BF_CUDA_OK;
}

BfCudaStatus validate(const BfCudaHelm2Build *b) {
  if (!b) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null build");
  if (b->numPairs < 0 || b->numRecipes < 0 || b->numOps < 0) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "negative count");
  if ((b->numPairs && !b->pairs) || (b->numRecipes && !b->recipes)) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null table");
  return BF_CUDA_OK;
}

} // namespace

extern "C" BfCudaStatus bfCudaPlanBuildHelm2(BfCudaPlan *P, const BfCudaHelm2Build *b, BfCudaHelm2Stats *stats) {
  if (!P) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "null plan");
  if (P->loadedFromFile) return fail(BF_CUDA_ERROR_NOT_IMPLEMENTED, "a plan loaded from a file holds no per-leaf data: its values cannot be re-built");
  BfCudaStatus st = validate(b);
  if (st != BF_CUDA_OK) return st;
  if (P->dtype != BF_CUDA_COMPLEX_F64) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "Helmholtz factors are complex");
  if (b->numOps != P->numOps || (b->numOps && !b->opRecipe)) return fail(BF_CUDA_ERROR_INCOMPATIBLE_SHAPES, "opRecipe does not match the plan's raw ops");
  if (stats) std::memset(stats, 0, sizeof *stats);
  DeviceGuard g(P->device);

  // which recipes are referenced, and do the leaves have the recipes' shapes
  std::vector<char> used((size_t)b->numRecipes, 0);
  const DevDirection *any = nullptr;
  for (const auto &dd : P->dir) if (dd.built) { any = &dd; break; }
  if (!any) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "plan has no built direction");
  for (const DirOp &d : any->host.ops) {
    const int64_t r = b->opRecipe[d.rawIndex];
    if (r < 0 || d.remote) continue; // block-row sharding: another rank builds (and applies) that leaf
    if (r >= b->numRecipes) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "opRecipe out of range");
    if (d.kind != BF_CUDA_LEAF_DENSE) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "recipe on a non-dense leaf");
    const BfCudaHelm2Recipe &c = b->recipes[r];
    const int64_t total = d.m * d.k;
    const int64_t rowOff = b->opRowOffset ? b->opRowOffset[d.rawIndex] : 0;
    if (d.storedCols != c.cols || c.cols <= 0 || total % c.cols != 0 || rowOff < 0 || rowOff + total / c.cols > c.rows ||
        (!b->opRowOffset && total != c.rows * c.cols))
      return fail(BF_CUDA_ERROR_INCOMPATIBLE_SHAPES, "leaf shape differs from its recipe");
    if (!used[(size_t)r] && stats) {
      if (c.kind == BF_CUDA_HELM2_KERNEL) { ++stats->kernelRecipes; stats->kernelEntries += total; }
      else { ++stats->reexpRecipes; stats->reexpEntries += total; }
    }
    used[(size_t)r] = 1;
  }

  cudaDeviceProp prop;
  BF_CUDA_TRY(cudaGetDeviceProperties(&prop, P->device));
  Prepared prep;
  st = prepare(*b, used, prep, (size_t)prop.sharedMemPerBlockOptin - 2048);
  if (st != BF_CUDA_OK) return st;
  DeviceBuffers mem;
  Uploaded u;
  st = runPinv(*b, prep, mem, u, P->numSMs, P->stream, stats);
  if (st != BF_CUDA_OK) return st;

  BF_CUDA_TRY(cudaFuncSetAttribute(bfHelm2Fill, cudaFuncAttributeMaxDynamicSharedMemorySize, kFillSmemBytes));
  BF_CUDA_TRY(cudaFuncSetAttribute(bfHelm2Fill, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  // every tile of every stage of every built direction, in ONE launch (tiles are independent)
  std::vector<FillTile> tiles;
  for (auto &dd : P->dir) {
    if (!dd.built) continue;
    const Direction &D = dd.host;
    for (const Stage &S : D.stages) {
      for (size_t t = 0; t < S.tiles.size(); ++t) {
        if (S.tiles[t].kc == kKcIdentity || S.tiles[t].kc == kKcDiag) continue;
        const DirOp &d = D.ops[(size_t)S.pieces[t].op];
        const int64_t r = b->opRecipe[d.rawIndex];
        if (r < 0) continue;
        FillTile f;
        f.dst = (uint64_t)(uintptr_t)(dd.pool + S.tilePool16[t] * 16ull);
        f.recipe = (int32_t)r;
        f.row0 = S.pieces[t].row0; f.col0 = S.pieces[t].col0;
        // a leaf that is a row slab of its recipe: shift the STORED row (acting column when the direction swaps)
        if (b->opRowOffset) (d.swapRC ? f.col0 : f.row0) += (int32_t)b->opRowOffset[d.rawIndex];
        f.rows = S.tiles[t].rows; f.kc = S.tiles[t].kc;
        f.swapRC = d.swapRC; f.conj = d.conj; f.pad = 0;
        tiles.push_back(f);
      }
    }
  }
  float ms = 0;
  st = runFill(tiles, u, mem, P->numSMs, P->stream, &ms);
  if (st != BF_CUDA_OK) return st;
  if (stats) stats->packMs = ms;
  return BF_CUDA_OK;
}

// One recipe on its own: the matrix is produced tile by tile exactly as for a plan (32 x 16 tiles into a
// scratch "pool") and re-assembled row-major on the host.
extern "C" BfCudaStatus bfCudaHelm2Evaluate(const BfCudaHelm2Build *b, int64_t index, void *out, int32_t device) {
  BfCudaStatus st = validate(b);
  if (st != BF_CUDA_OK) return st;
  if (index < 0 || index >= b->numRecipes || !out) return fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "recipe index");
  if (device < 0) BF_CUDA_TRY(cudaGetDevice(&device));
  DeviceGuard g(device);
  cudaDeviceProp prop;
  BF_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) return fail(BF_CUDA_ERROR_NO_DEVICE, "an sm_100-class device is required");
  std::vector<char> used((size_t)b->numRecipes, 0);
  used[(size_t)index] = 1;
  Prepared prep;
  st = prepare(*b, used, prep, (size_t)prop.sharedMemPerBlockOptin - 2048);
  if (st != BF_CUDA_OK) return st;
  DeviceBuffers mem;
  Uploaded u;
  cudaStream_t stream = nullptr;
  st = runPinv(*b, prep, mem, u, prop.multiProcessorCount, stream, nullptr);
  if (st != BF_CUDA_OK) return st;
  const BfCudaHelm2Recipe &c = b->recipes[index];
  const int TR = 32, TC = 16;
  std::vector<FillTile> tiles;
  uint64_t pool16 = 0;
  for (int64_t r0 = 0; r0 < c.rows; r0 += TR)
    for (int64_t c0 = 0; c0 < c.cols; c0 += TC) {
      FillTile f;
      std::memset(&f, 0, sizeof f);
      f.dst = pool16; f.recipe = (int32_t)index; // offset in elements for now, rebased below
      f.row0 = (int32_t)r0; f.col0 = (int32_t)c0;
      f.rows = (uint16_t)std::min<int64_t>(TR, c.rows - r0); f.kc = (uint16_t)std::min<int64_t>(TC, c.cols - c0);
      pool16 += (uint64_t)f.rows * f.kc;
      tiles.push_back(f);
    }
  unsigned char *pool = nullptr;
  BF_CUDA_TRY(mem.alloc(&pool, (size_t)pool16 * 16 + 16));
  std::vector<uint64_t> offs(tiles.size());
  for (size_t t = 0; t < tiles.size(); ++t) { offs[t] = tiles[t].dst; tiles[t].dst = (uint64_t)(uintptr_t)(pool + offs[t] * 16ull); }
  BF_CUDA_TRY(cudaFuncSetAttribute(bfHelm2Fill, cudaFuncAttributeMaxDynamicSharedMemorySize, kFillSmemBytes));
  st = runFill(tiles, u, mem, prop.multiProcessorCount, stream, nullptr);
  if (st != BF_CUDA_OK) return st;
  std::vector<double2> host((size_t)pool16);
  if (pool16) BF_CUDA_TRY(cudaMemcpy(host.data(), pool, (size_t)pool16 * 16, cudaMemcpyDeviceToHost));
  double2 *dst = static_cast<double2 *>(out);
  for (size_t t = 0; t < tiles.size(); ++t) {
    const FillTile &f = tiles[t];
    for (int r = 0; r < f.rows; ++r)
      for (int q = 0; q < f.kc; ++q)
        dst[(size_t)(f.row0 + r) * (size_t)c.cols + (size_t)(f.col0 + q)] = host[(size_t)offs[t] + (size_t)r * f.kc + q];
  }
  return BF_CUDA_OK;
}
