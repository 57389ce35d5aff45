// bf_peak.cu — measured fp64 ceilings of the device the plan runs on: the denominators of the
// many-RHS roofline (MEASURED_PEAKS.json carries HBM and bf16 figures only).  Register-resident
// loops, no memory traffic: (1) DFMA, (2) DMMA mma.sync.m8n8k4.f64.
#include <cuda_runtime.h>

#include <algorithm>

#include "bf_cuda.h"

namespace {
template <int ILP>
__global__ void bfPeakDfma(double *out, int iters, double a, double b) {
  double acc[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) acc[i] = threadIdx.x + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void bfPeakDmma(double *out, int iters, double a, double b) {
  double c0[ILP], c1[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) { c0[i] = threadIdx.x + i; c1[i] = i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c0[i]), "+d"(c1[i]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += c0[i] + c1[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
} // namespace

extern "C" BfCudaStatus bfCudaMeasureFp64Peak(int32_t device, double *dfmaTflops, double *dmmaTflops) {
  int prev = -1, nDev = 0;
  if (cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0) { cudaGetLastError(); return BF_CUDA_ERROR_NO_DEVICE; }
  cudaGetDevice(&prev);
  if (device < 0) device = prev;
  if (device >= nDev || cudaSetDevice(device) != cudaSuccess) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, device);
  const int sms = p.multiProcessorCount, threads = 256, ctas = sms * 2, iters = 20000;
  double *out = nullptr;
  cudaEvent_t e0, e1;
  BfCudaStatus st = BF_CUDA_OK;
  if (cudaMalloc(&out, sizeof(double) * ctas * threads) != cudaSuccess) { cudaSetDevice(prev); return BF_CUDA_ERROR_OUT_OF_MEMORY; }
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best[2] = {0.0, 0.0};
  for (int which = 0; which < 2; ++which)
    for (int rep = 0; rep < 4; ++rep) { // rep 0 warms up
      cudaEventRecord(e0);
      if (which == 0) bfPeakDfma<16><<<ctas, threads>>>(out, iters, 1.0000001, 1e-9);
      else bfPeakDmma<8><<<ctas, threads>>>(out, iters, 1.0000001, 1e-9);
      cudaEventRecord(e1);
      if (cudaEventSynchronize(e1) != cudaSuccess) { st = BF_CUDA_ERROR_RUNTIME; break; }
      float ms = 0.f;
      cudaEventElapsedTime(&ms, e0, e1);
      // DFMA: 2 flop x 16 per thread-iteration; one m8n8k4 = 256 FMA = 512 flop per warp
      const double flops = which == 0 ? 2.0 * 16 * iters * (double)threads * ctas
                                      : 512.0 * 8 * iters * (threads / 32.0) * ctas;
      if (rep) best[which] = std::max(best[which], flops / ms / 1e9);
    }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(out);
  if (prev >= 0) cudaSetDevice(prev);
  if (dfmaTflops) *dfmaTflops = best[0];
  if (dmmaTflops) *dmmaTflops = best[1];
  return st;
}

// Sustained variant: the DMMA loop launched back to back for `seconds`; the figure is the average over
// the second half, i.e. after the power management has settled (a 1 kW part drops its SM clock under
// a dense fp64 tensor load).  This is the denominator for a kernel timed inside a long step.
extern "C" BfCudaStatus bfCudaMeasureFp64PeakSustained(int32_t device, double seconds, double *dmmaTflops) {
  int prev = -1, nDev = 0;
  if (cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0) { cudaGetLastError(); return BF_CUDA_ERROR_NO_DEVICE; }
  cudaGetDevice(&prev);
  if (device < 0) device = prev;
  if (device >= nDev || cudaSetDevice(device) != cudaSuccess) return BF_CUDA_ERROR_INVALID_ARGUMENT;
  if (!(seconds > 0.0) || seconds > 30.0) seconds = 2.0;
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, device);
  const int sms = p.multiProcessorCount, threads = 256, ctas = sms * 2, iters = 20000;
  double *out = nullptr;
  if (cudaMalloc(&out, sizeof(double) * ctas * threads) != cudaSuccess) { cudaSetDevice(prev); return BF_CUDA_ERROR_OUT_OF_MEMORY; }
  cudaEvent_t e0, e1, e2;
  cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
  // one launch is ~5 ms at full clock
  bfPeakDmma<8><<<ctas, threads>>>(out, iters, 1.0000001, 1e-9);
  cudaDeviceSynchronize();
  const int launches = (int)(seconds / 0.0055) + 2, half = launches / 2;
  cudaEventRecord(e0);
  for (int i = 0; i < launches; ++i) {
    if (i == half) cudaEventRecord(e1);
    bfPeakDmma<8><<<ctas, threads>>>(out, iters, 1.0000001, 1e-9);
  }
  cudaEventRecord(e2);
  BfCudaStatus st = cudaEventSynchronize(e2) == cudaSuccess ? BF_CUDA_OK : BF_CUDA_ERROR_RUNTIME;
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e1, e2);
  const double flops = 512.0 * 8 * iters * (threads / 32.0) * ctas * (double)(launches - half);
  if (dmmaTflops) *dmmaTflops = st == BF_CUDA_OK ? flops / ms / 1e9 : 0.0;
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
  cudaFree(out);
  if (prev >= 0) cudaSetDevice(prev);
  return st;
}
