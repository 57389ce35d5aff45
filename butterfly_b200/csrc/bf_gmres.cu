// bf_gmres.cu — device-resident GMRES around the butterfly apply.
//
// Restates bfSolveGMRES (src/linalg.c:47-318) for one complex right-hand side and no preconditioner:
// modified Gram-Schmidt Arnoldi (src/linalg.c:155-190), Givens rotations computed the way
// bfVecComplexGetGivensRotation does (src/vec_complex.c:230-290) and applied as
// mulInplace_givensComplex does (src/vec_complex.c:155-168), the convergence test
// |s[j+1]| / beta < tol (src/linalg.c:224-231) and the reconstruction from the first j basis vectors
// (src/linalg.c:237-281 — when the test fires at column j the reference leaves the loop BEFORE
// counting that column, so the returned iterate uses j columns; we do the same).
//
// The reference round-trips every Krylov vector through host malloc and does the Gram-Schmidt with
// scalar loops.  Here the basis V, the work vector and all dots stay in HBM; each Gram-Schmidt step
// is ONE kernel (reduce the previous step's partial dots -> h_i, w -= h_i v_i, partial dots against
// v_{i+1}), reductions are two-stage and fixed-order (deterministic), and only the j + 2 entries of
// the new Hessenberg column cross PCIe per iteration.
#include <cuda_runtime.h>

#include <cmath>
#include <complex>
#include <mutex>
#include <string>
#include <vector>

#include "bf_cuda.h"

namespace bfc {
BfCudaStatus fail(BfCudaStatus st, const std::string &msg); // bf_capi.cu: records bfCudaLastErrorMessage
}

namespace {
constexpr int kBlocks = 148 * 4;
constexpr int kThreads = 256;

__device__ __forceinline__ double2 blockSum(double2 v, double2 *sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
    v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  double2 t = make_double2(0.0, 0.0);
  for (int i = 0; i < kThreads / 32; ++i) { t.x += sh[i].x; t.y += sh[i].y; } // same order in every thread
  return t;
}
// Total of `num` partials, identical in every thread of every block (fixed order).
__device__ __forceinline__ double2 totalOf(const double2 *partials, int num, double2 *sh) {
  double2 v = make_double2(0.0, 0.0);
  for (int i = threadIdx.x; i < num; i += kThreads) { v.x += partials[i].x; v.y += partials[i].y; }
  return blockSum(v, sh);
}

// partials[b] = sum over the block's elements of conj(v) * w   (zdotc, src/mat_dense_complex.c:115-121)
// or of |w|^2 when v == NULL.
__device__ __forceinline__ void dotInto(const double2 *v, const double2 *w, long long n, double2 *partials, double2 *sh) {
  double2 acc = make_double2(0.0, 0.0);
  for (long long k = (long long)blockIdx.x * kThreads + threadIdx.x; k < n; k += (long long)gridDim.x * kThreads) {
    const double2 b = w[k];
    if (v) {
      const double2 a = v[k];
      acc.x += a.x * b.x + a.y * b.y;
      acc.y += a.x * b.y - a.y * b.x;
    } else {
      acc.x += b.x * b.x + b.y * b.y;
    }
  }
  acc = blockSum(acc, sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
}

__global__ void __launch_bounds__(kThreads) gmresDot(const double2 *v, const double2 *w, long long n, double2 *partials) {
  __shared__ double2 sh[kThreads / 32];
  dotInto(v, w, n, partials, sh);
}

// h = total(partialsIn); hOut = h; w -= h * vi; partialsOut = <vNext, w> (or |w|^2 when vNext == NULL)
__global__ void __launch_bounds__(kThreads) gmresOrthoStep(const double2 *partialsIn, int numPartials, const double2 *vi,
                                                            double2 *w, const double2 *vNext, long long n,
                                                            double2 *partialsOut, double2 *hOut) {
  __shared__ double2 sh[kThreads / 32];
  const double2 h = totalOf(partialsIn, numPartials, sh);
  if (blockIdx.x == 0 && threadIdx.x == 0) *hOut = h;
  double2 acc = make_double2(0.0, 0.0);
  for (long long k = (long long)blockIdx.x * kThreads + threadIdx.x; k < n; k += (long long)gridDim.x * kThreads) {
    const double2 a = vi[k];
    double2 b = w[k];
    b.x -= h.x * a.x - h.y * a.y;
    b.y -= h.x * a.y + h.y * a.x;
    w[k] = b;
    if (vNext) {
      const double2 c = vNext[k];
      acc.x += c.x * b.x + c.y * b.y;
      acc.y += c.x * b.y - c.y * b.x;
    } else {
      acc.x += b.x * b.x + b.y * b.y;
    }
  }
  acc = blockSum(acc, sh);
  if (threadIdx.x == 0) partialsOut[blockIdx.x] = acc;
}

// nrm = sqrt(total(partials).x); normOut = nrm; out = (src - sub) / nrm   (sub may be NULL)
__global__ void __launch_bounds__(kThreads) gmresNormalize(const double2 *partials, int numPartials, const double2 *src,
                                                            const double2 *sub, long long n, double2 *out, double2 *normOut) {
  __shared__ double2 sh[kThreads / 32];
  const double nrm = sqrt(totalOf(partials, numPartials, sh).x);
  if (blockIdx.x == 0 && threadIdx.x == 0) *normOut = make_double2(nrm, 0.0);
  for (long long k = (long long)blockIdx.x * kThreads + threadIdx.x; k < n; k += (long long)gridDim.x * kThreads) {
    double2 a = src[k];
    out[k] = make_double2(a.x / nrm, a.y / nrm);
  }
}

__global__ void gmresSub(const double2 *a, const double2 *b, long long n, double2 *out) { // out = a - b
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x)
    out[k] = make_double2(a[k].x - b[k].x, a[k].y - b[k].y);
}

// x = x0 + sum_{i<m} y[i] * V[i]   (src/linalg.c:268-272)
__global__ void gmresCombine(const double2 *V, long long n, const double2 *y, int m, const double2 *x0, double2 *x) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
    double2 acc = make_double2(0.0, 0.0);
    for (int i = 0; i < m; ++i) {
      const double2 v = V[(long long)i * n + k], c = y[i];
      acc.x += c.x * v.x - c.y * v.y;
      acc.y += c.x * v.y + c.y * v.x;
    }
    if (x0) { acc.x += x0[k].x; acc.y += x0[k].y; }
    x[k] = acc;
  }
}

// ---- building blocks of the ROW-DISTRIBUTED solve (butterfly_b200/gmres.py) -------------------------
// Every rank holds its rows of the Krylov basis; a Gram-Schmidt pass is "local partial sums -> one
// all_reduce -> local update".  These kernels are the local halves: all projections of a pass in ONE
// sweep over the work vector, fixed-order two-stage sums (bitwise reproducible run to run).
constexpr int kProjChunk = 8;

// partials[(c) * gridDim.x + b] = sum over block b's elements of conj(V[c]) * w, c < mc;
// if normSlot >= 0 also partials[normSlot * gridDim.x + b] = sum |w|^2.
__global__ void __launch_bounds__(kThreads) gmresProjectPartial(const double2 *__restrict__ V, long long ldv, int mc,
                                                                 const double2 *__restrict__ w, long long n,
                                                                 double2 *__restrict__ partials, int normSlot) {
  __shared__ double2 sh[kThreads / 32];
  double2 acc[kProjChunk];
  double nrm = 0.0;
#pragma unroll
  for (int c = 0; c < kProjChunk; ++c) acc[c] = make_double2(0.0, 0.0);
  for (long long k = (long long)blockIdx.x * kThreads + threadIdx.x; k < n; k += (long long)gridDim.x * kThreads) {
    const double2 b = w[k];
    nrm += b.x * b.x + b.y * b.y;
#pragma unroll
    for (int c = 0; c < kProjChunk; ++c)
      if (c < mc) {
        const double2 a = V[(long long)c * ldv + k];
        acc[c].x += a.x * b.x + a.y * b.y;
        acc[c].y += a.x * b.y - a.y * b.x;
      }
  }
#pragma unroll
  for (int c = 0; c < kProjChunk; ++c)
    if (c < mc) {
      const double2 t = blockSum(acc[c], sh);
      if (threadIdx.x == 0) partials[(long long)c * gridDim.x + blockIdx.x] = t;
    }
  if (normSlot >= 0) {
    const double2 t = blockSum(make_double2(nrm, 0.0), sh);
    if (threadIdx.x == 0) partials[(long long)normSlot * gridDim.x + blockIdx.x] = t;
  }
}

// out[i] = sum_b partials[i * num + b], one block per output, fixed order
__global__ void __launch_bounds__(kThreads) gmresFinalize(const double2 *__restrict__ partials, int num, double2 *__restrict__ out) {
  __shared__ double2 sh[kThreads / 32];
  const double2 t = totalOf(partials + (long long)blockIdx.x * num, num, sh);
  if (threadIdx.x == 0) out[blockIdx.x] = t;
}

// w -= sum_{i<m} h[i] V[i];  partials[b] = sum |w_new|^2 over block b
__global__ void __launch_bounds__(kThreads) gmresSubtractPartial(const double2 *__restrict__ V, long long ldv, int m,
                                                                  const double2 *__restrict__ h, double2 *__restrict__ w,
                                                                  long long n, double2 *__restrict__ partials) {
  __shared__ double2 sh[kThreads / 32];
  extern __shared__ double2 hs[];
  for (int i = threadIdx.x; i < m; i += kThreads) hs[i] = h[i];
  __syncthreads();
  double nrm = 0.0;
  for (long long k = (long long)blockIdx.x * kThreads + threadIdx.x; k < n; k += (long long)gridDim.x * kThreads) {
    double2 b = w[k];
#pragma unroll 4
    for (int i = 0; i < m; ++i) {
      const double2 a = V[(long long)i * ldv + k], c = hs[i];
      b.x -= c.x * a.x - c.y * a.y;
      b.y -= c.x * a.y + c.y * a.x;
    }
    w[k] = b;
    nrm += b.x * b.x + b.y * b.y;
  }
  const double2 t = blockSum(make_double2(nrm, 0.0), sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

// dst = w / sqrt(nrm2->x)  (and dst2, when given: the all_gather staging copy)
__global__ void __launch_bounds__(kThreads) gmresScale(const double2 *__restrict__ w, long long n, const double2 *__restrict__ nrm2,
                                                        double2 *__restrict__ dst, double2 *__restrict__ dst2) {
  const double nrm = sqrt(nrm2->x);
  for (long long k = (long long)blockIdx.x * kThreads + threadIdx.x; k < n; k += (long long)gridDim.x * kThreads) {
    const double2 a = w[k];
    const double2 r = make_double2(a.x / nrm, a.y / nrm);
    dst[k] = r;
    if (dst2) dst2[k] = r;
  }
}

using cplx = std::complex<double>;
struct Givens { cplx c, s; };
// src/vec_complex.c:276-287
Givens makeGivens(cplx a, cplx b) {
  Givens g;
  if (std::abs(b) == 0) { g.c = 1; g.s = 0; }
  else if (std::abs(b) > std::abs(a)) { cplx t = -a / b; g.s = 1.0 / std::sqrt(1 + std::pow(std::abs(t), 2)); g.c = t * g.s; }
  else { cplx t = -b / a; g.c = 1.0 / std::sqrt(1 + std::pow(std::abs(t), 2)); g.s = t * g.c; }
  return g;
}
// src/vec_complex.c:155-168
void applyGivens(const Givens &g, cplx &z0, cplx &z1) {
  const cplx a = z0, b = z1;
  z0 = std::conj(g.c) * a + -g.s * b;
  z1 = g.s * a + g.c * b;
}
} // namespace

#define GM_TRY(expr)                                                                              \
  do {                                                                                            \
    cudaError_t e_ = (expr);                                                                      \
    if (e_ != cudaSuccess) { st = e_ == cudaErrorMemoryAllocation ? BF_CUDA_ERROR_OUT_OF_MEMORY : BF_CUDA_ERROR_RUNTIME; goto done; } \
  } while (0)

static BfCudaStatus solveImpl(BfCudaPlan *plan, const void *b, const void *x0, double tol, int64_t maxNumIter, void *x,
                              int64_t *numIter, double *relResidual, int64_t *numApplies, BfCudaPrecondFn precond, void *precondCtx) {
  if (!plan || !b || !x || maxNumIter <= 0) return BF_CUDA_ERROR_INVALID_ARGUMENT; // src/linalg.c:80-81
  BfCudaPlanInfo info;
  BfCudaStatus st = bfCudaPlanGetInfo(plan, &info);
  if (st != BF_CUDA_OK) return st;
  if (info.dtype != BF_CUDA_COMPLEX_F64) return BF_CUDA_ERROR_NOT_IMPLEMENTED;
  if (info.numRows != info.numCols) return BF_CUDA_ERROR_INVALID_ARGUMENT; // src/linalg.c:85-86
  const long long n = info.numRows;
  const int m = (int)maxNumIter;
  int prevDev = -1;
  cudaGetDevice(&prevDev);
  if (prevDev != info.device) cudaSetDevice(info.device);
  cudaStream_t stream = (cudaStream_t)bfCudaPlanGetStream(plan);
  double2 *V = nullptr, *w = nullptr, *partials = nullptr, *hcol = nullptr, *dx0 = nullptr, *dy = nullptr;
  std::vector<cplx> S((size_t)m + 1, cplx(0)), hhost((size_t)m + 2), Hall((size_t)m * ((size_t)m + 1), cplx(0));
  std::vector<Givens> J;
  std::vector<cplx> hostIn, hostOut; // left preconditioner: the vector crosses PCIe once each way
  if (precond) { hostIn.resize((size_t)n); hostOut.resize((size_t)n); }
  // w <- M^{-1} w through the caller's host routine (the reference: bfMatSolve(M, W), src/linalg.c:131,159)
  auto applyPrecond = [&](double2 *w) -> BfCudaStatus {
    if (cudaMemcpyAsync(hostIn.data(), w, (size_t)n * sizeof(double2), cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
        cudaStreamSynchronize(stream) != cudaSuccess)
      return BF_CUDA_ERROR_RUNTIME;
    if (precond(precondCtx, hostIn.data(), hostOut.data(), n) != 0) return BF_CUDA_ERROR_RUNTIME;
    if (cudaMemcpyAsync(w, hostOut.data(), (size_t)n * sizeof(double2), cudaMemcpyHostToDevice, stream) != cudaSuccess)
      return BF_CUDA_ERROR_RUNTIME;
    return BF_CUDA_OK;
  };
  double beta = 0.0, residual = 0.0;
  int64_t applies = 0;
  int j = 0;
  bool converged = false;
  const size_t vecBytes = (size_t)n * sizeof(double2);
  GM_TRY(cudaMalloc(&V, ((size_t)m + 1) * vecBytes));
  GM_TRY(cudaMalloc(&w, vecBytes));
  GM_TRY(cudaMalloc(&partials, 2 * kBlocks * sizeof(double2)));
  GM_TRY(cudaMalloc(&hcol, ((size_t)m + 2) * sizeof(double2)));
  GM_TRY(cudaMalloc(&dy, ((size_t)m + 1) * sizeof(double2)));
  // R = B - A X0 (src/linalg.c:124-129); X0 == NULL is the zero iterate, for which A X0 == 0 exactly
  GM_TRY(cudaMemcpyAsync(w, b, vecBytes, cudaMemcpyHostToDevice, stream));
  if (x0) {
    GM_TRY(cudaMalloc(&dx0, vecBytes));
    GM_TRY(cudaMemcpyAsync(dx0, x0, vecBytes, cudaMemcpyHostToDevice, stream));
    st = bfCudaApplyDevice(plan, BF_CUDA_OP_N, dx0, 1, 1, V, 1, stream); // V[0] as scratch
    if (st != BF_CUDA_OK) goto done;
    ++applies;
    gmresSub<<<kBlocks, kThreads, 0, stream>>>(w, V, n, w);
  }
  if (precond) { st = applyPrecond(w); if (st != BF_CUDA_OK) goto done; }
  // beta = ||r||, V[0] = r / beta, S = beta e1 (src/linalg.c:136-148)
  gmresDot<<<kBlocks, kThreads, 0, stream>>>(nullptr, w, n, partials);
  gmresNormalize<<<kBlocks, kThreads, 0, stream>>>(partials, kBlocks, w, nullptr, n, V, hcol);
  GM_TRY(cudaMemcpyAsync(hhost.data(), hcol, sizeof(double2), cudaMemcpyDeviceToHost, stream));
  GM_TRY(cudaStreamSynchronize(stream));
  beta = hhost[0].real();
  S[0] = beta;
  if (!(beta > 0.0)) { // zero residual: the initial iterate is the solution
    converged = true;
  }
  for (j = 0; j < m && !converged; ++j) {
    double2 *Vj = V + (size_t)j * n;
    st = bfCudaApplyDevice(plan, BF_CUDA_OP_N, Vj, 1, 1, w, 1, stream); // W = A V[j] (src/linalg.c:155)
    if (st != BF_CUDA_OK) goto done;
    ++applies;
    if (precond) { st = applyPrecond(w); if (st != BF_CUDA_OK) goto done; }
    // modified Gram-Schmidt (src/linalg.c:172-186), one kernel per basis vector
    double2 *pIn = partials, *pOut = partials + kBlocks;
    gmresDot<<<kBlocks, kThreads, 0, stream>>>(V, w, n, pIn);
    for (int i = 0; i <= j; ++i) {
      const double2 *vNext = i < j ? V + (size_t)(i + 1) * n : nullptr;
      gmresOrthoStep<<<kBlocks, kThreads, 0, stream>>>(pIn, kBlocks, V + (size_t)i * n, w, vNext, n, pOut, hcol + i);
      std::swap(pIn, pOut);
    }
    // H[j+1, j] = ||w||, V[j+1] = w / ||w|| (src/linalg.c:188-202)
    gmresNormalize<<<kBlocks, kThreads, 0, stream>>>(pIn, kBlocks, w, nullptr, n, V + (size_t)(j + 1) * n, hcol + j + 1);
    GM_TRY(cudaGetLastError());
    GM_TRY(cudaMemcpyAsync(hhost.data(), hcol, ((size_t)j + 2) * sizeof(double2), cudaMemcpyDeviceToHost, stream));
    GM_TRY(cudaStreamSynchronize(stream));
    // previous rotations, the new one, the rotated right-hand side (src/linalg.c:206-231)
    for (int i = 0; i < j; ++i) applyGivens(J[(size_t)i], hhost[(size_t)i], hhost[(size_t)i + 1]);
    J.push_back(makeGivens(hhost[(size_t)j], hhost[(size_t)j + 1]));
    applyGivens(J.back(), hhost[(size_t)j], hhost[(size_t)j + 1]);
    applyGivens(J.back(), S[(size_t)j], S[(size_t)j + 1]);
    for (int i = 0; i <= j; ++i) Hall[(size_t)j * ((size_t)m + 1) + i] = hhost[(size_t)i];
    residual = std::abs(S[(size_t)j + 1]) / beta;
    if (residual < tol) { converged = true; break; } // leaves j uncounted, as the reference does
  }
  {
    // back substitution on the leading j x j triangle (src/linalg.c:240-266), x = x0 + V[:, :j] y
    std::vector<cplx> y((size_t)std::max(j, 1), cplx(0));
    for (int r = j - 1; r >= 0; --r) {
      cplx acc = S[(size_t)r];
      for (int c = r + 1; c < j; ++c) acc -= Hall[(size_t)c * ((size_t)m + 1) + r] * y[(size_t)c];
      y[(size_t)r] = acc / Hall[(size_t)r * ((size_t)m + 1) + r];
    }
    if (j > 0) GM_TRY(cudaMemcpyAsync(dy, y.data(), (size_t)j * sizeof(double2), cudaMemcpyHostToDevice, stream));
    gmresCombine<<<kBlocks, kThreads, 0, stream>>>(V, n, dy, j, dx0, w);
    GM_TRY(cudaGetLastError());
    GM_TRY(cudaMemcpyAsync(x, w, vecBytes, cudaMemcpyDeviceToHost, stream));
    GM_TRY(cudaStreamSynchronize(stream));
  }
  if (numIter) *numIter = j;
  if (relResidual) *relResidual = residual;
  if (numApplies) *numApplies = applies;
  st = BF_CUDA_OK;
done:
  cudaFree(V); cudaFree(w); cudaFree(partials); cudaFree(hcol); cudaFree(dx0); cudaFree(dy);
  if (prevDev >= 0 && prevDev != info.device) cudaSetDevice(prevDev);
  return st;
}

extern "C" BfCudaStatus bfCudaSolveGMRES(BfCudaPlan *plan, const void *b, const void *x0, double tol, int64_t maxNumIter,
                                         void *x, int64_t *numIter, double *relResidual, int64_t *numApplies) {
  return solveImpl(plan, b, x0, tol, maxNumIter, x, numIter, relResidual, numApplies, nullptr, nullptr);
}

extern "C" BfCudaStatus bfCudaSolveGMRESPrecond(BfCudaPlan *plan, const void *b, const void *x0, double tol,
                                                int64_t maxNumIter, void *x, int64_t *numIter, double *relResidual,
                                                int64_t *numApplies, BfCudaPrecondFn precond, void *precondCtx) {
  return solveImpl(plan, b, x0, tol, maxNumIter, x, numIter, relResidual, numApplies, precond, precondCtx);
}

// ---- C ABI of the row-distributed building blocks (device pointers, caller's stream) ------------------
extern "C" int64_t bfCudaGmresScratchBytes(int64_t m) { return (int64_t)sizeof(double2) * kBlocks * (m + 2); }

extern "C" BfCudaStatus bfCudaGmresProject(const void *V, int64_t ldv, int64_t m, const void *w, int64_t n, void *out,
                                           void *scratch, void *stream) {
  if (m < 0 || n < 0 || !w || !out || !scratch || (m > 0 && !V)) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  cudaStream_t s = (cudaStream_t)stream;
  double2 *part = (double2 *)scratch;
  const double2 *Vd = (const double2 *)V;
  // chunks of kProjChunk basis vectors per sweep; the last sweep also sums |w|^2 into slot m
  for (int64_t i0 = 0; i0 == 0 || i0 < m; i0 += kProjChunk) {
    const int mc = (int)std::min<int64_t>(kProjChunk, m - i0);
    const bool last = i0 + kProjChunk >= m;
    gmresProjectPartial<<<kBlocks, kThreads, 0, s>>>(Vd ? Vd + i0 * ldv : nullptr, ldv, mc, (const double2 *)w, n,
                                                     part + i0 * kBlocks, last ? mc : -1);
  }
  gmresFinalize<<<(unsigned)(m + 1), kThreads, 0, s>>>(part, kBlocks, (double2 *)out);
  return cudaGetLastError() == cudaSuccess ? BF_CUDA_OK : BF_CUDA_ERROR_RUNTIME;
}

extern "C" BfCudaStatus bfCudaGmresSubtract(const void *V, int64_t ldv, int64_t m, const void *h, void *w, int64_t n,
                                            void *nrm2Out, void *scratch, void *stream) {
  if (m < 0 || n < 0 || !w || !nrm2Out || !scratch || (m > 0 && (!V || !h))) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  cudaStream_t s = (cudaStream_t)stream;
  gmresSubtractPartial<<<kBlocks, kThreads, (size_t)std::max<int64_t>(m, 1) * sizeof(double2), s>>>(
      (const double2 *)V, ldv, (int)m, (const double2 *)h, (double2 *)w, n, (double2 *)scratch);
  gmresFinalize<<<1, kThreads, 0, s>>>((const double2 *)scratch, kBlocks, (double2 *)nrm2Out);
  return cudaGetLastError() == cudaSuccess ? BF_CUDA_OK : BF_CUDA_ERROR_RUNTIME;
}

extern "C" BfCudaStatus bfCudaGmresScale(const void *w, int64_t n, const void *nrm2, void *dst, void *dst2, void *stream) {
  if (n < 0 || !w || !nrm2 || !dst) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  gmresScale<<<kBlocks, kThreads, 0, (cudaStream_t)stream>>>((const double2 *)w, n, (const double2 *)nrm2, (double2 *)dst, (double2 *)dst2);
  return cudaGetLastError() == cudaSuccess ? BF_CUDA_OK : BF_CUDA_ERROR_RUNTIME;
}

// ---- allocation-free variant of bfCudaGmresProject ------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The partial sums
// live in a file-scope device array instead of a caller-provided scratch buffer: kBlocks x (maxM + 2) double2 =
// 606 208 bytes, held by every GPU that loads the library until the process exits.
//
// In HOST code the array's name is a host-side stand-in, not a device address (a kernel handed it faults, or —
// where the GPU can reach host memory — silently uses the wrong memory), so the address is asked for once per
// device with that device current (cudaGetSymbolAddress, return code checked) and only that pointer is ever
// passed to a kernel.  The call path below contains no cudaMalloc / cudaFree / allocating library routine.
//
// All calls on a device SHARE the array: two calls that overlap on the GPU (different streams, no dependency
// between them) race on it — serialise them.  The array keeps the previous call's contents; nothing needs
// clearing, because every slot gmresFinalize reads (m + 1 rows of kBlocks partials) is written by the
// gmresProjectPartial launches of the SAME call before it, in stream order.
namespace {
constexpr int64_t kStaticProjectMaxM = BF_CUDA_GMRES_PROJECT_STATIC_MAX_M;
constexpr size_t kStaticScratchBytes = sizeof(double2) * (size_t)kBlocks * (size_t)(kStaticProjectMaxM + 2);
static __device__ __align__(256) unsigned char g_projectScratch[kStaticScratchBytes];

constexpr int kMaxScratchDevices = 64;
double2 *g_projectScratchAddr[kMaxScratchDevices]; // per device ordinal; NULL = not asked for yet
std::mutex g_projectScratchMutex;

// Device address of g_projectScratch on the CURRENT device (looked up once per device).
BfCudaStatus staticProjectScratch(double2 **ptr) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxScratchDevices)
    return bfc::fail(BF_CUDA_ERROR_NO_DEVICE, "bfCudaGmresProjectStatic: no current CUDA device");
  std::lock_guard<std::mutex> lock(g_projectScratchMutex);
  if (!g_projectScratchAddr[dev]) {
    double2 *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_projectScratch); // `dev` is current
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return bfc::fail(BF_CUDA_ERROR_RUNTIME, std::string("bfCudaGmresProjectStatic: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    }
    g_projectScratchAddr[dev] = p;
  }
  *ptr = g_projectScratchAddr[dev];
  return BF_CUDA_OK;
}
} // namespace

extern "C" BfCudaStatus bfCudaGmresProjectStatic(const void *V, int64_t ldv, int64_t m, const void *w, int64_t n, void *out,
                                                 void *stream) {
  if (m < 0 || n < 0 || !w || !out || (m > 0 && !V))
    return bfc::fail(BF_CUDA_ERROR_INVALID_ARGUMENT, "bfCudaGmresProjectStatic: null operand or negative size");
  if (m > kStaticProjectMaxM)
    return bfc::fail(BF_CUDA_ERROR_INVALID_ARGUMENT,
                     "bfCudaGmresProjectStatic: m = " + std::to_string(m) + " basis vectors exceed the static scratch (maximum " +
                         std::to_string(kStaticProjectMaxM) + "); use bfCudaGmresProject with a caller-provided scratch buffer");
  double2 *part = nullptr;
  const BfCudaStatus st = staticProjectScratch(&part);
  if (st != BF_CUDA_OK) return st;
  cudaStream_t s = (cudaStream_t)stream;
  const double2 *Vd = (const double2 *)V;
  // the same launches as bfCudaGmresProject: chunks of kProjChunk basis vectors per sweep, |w|^2 into slot m
  for (int64_t i0 = 0; i0 == 0 || i0 < m; i0 += kProjChunk) {
    const int mc = (int)std::min<int64_t>(kProjChunk, m - i0);
    const bool last = i0 + kProjChunk >= m;
    gmresProjectPartial<<<kBlocks, kThreads, 0, s>>>(Vd ? Vd + i0 * ldv : nullptr, ldv, mc, (const double2 *)w, n,
                                                     part + i0 * kBlocks, last ? mc : -1);
  }
  gmresFinalize<<<(unsigned)(m + 1), kThreads, 0, s>>>(part, kBlocks, (double2 *)out);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return bfc::fail(BF_CUDA_ERROR_RUNTIME, std::string("bfCudaGmresProjectStatic: launch: ") + cudaGetErrorString(e));
  return BF_CUDA_OK;
}
