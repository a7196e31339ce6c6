// peaks.cu -- tahoe_measure_peaks: measured read bandwidth of shared memory and L2 on the current GPU, the
// roofline denominators SURVEY 8(d) asks for and the driver's MEASURED_PEAKS.json does not hold (it has
// HBM and bf16 only).  bench.py calls it in the same run as the kernels it rates; `python -m
// tahoe_b200.peaks` prints the same numbers as one JSON line.  CUDA events, best of 5.
#include <cstdio>
#include <cuda_runtime.h>

#include "forest.h"

namespace {

// conflict-free LDS.32 stream: every lane reads bank = lane, 8 independent loads in flight
__global__ void __launch_bounds__(1024, 1) k_smem(unsigned* out, int iters) {
  extern __shared__ unsigned s[];
  for (int i = threadIdx.x; i < 48 * 1024; i += blockDim.x) s[i] = i;
  __syncthreads();
  unsigned acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
  const unsigned a = (unsigned)__cvta_generic_to_shared(s + threadIdx.x);   // lane i <-> bank i
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < 32; k += 4) {           // 32 independent volatile loads per iteration
      unsigned v0, v1, v2, v3;
      asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v0) : "r"(a + (k + 0) * 4096));
      asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v1) : "r"(a + (k + 1) * 4096));
      asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v2) : "r"(a + (k + 2) * 4096));
      asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v3) : "r"(a + (k + 3) * 4096));
      acc0 += v0; acc1 += v1; acc2 += v2; acc3 += v3;
    }
  }
  if ((acc0 ^ acc1 ^ acc2 ^ acc3) == 0xdeadbeef) out[0] = acc0;
}

// random-bank LDS.32: what a traversal's node loads look like (32 lanes on random words)
__global__ void __launch_bounds__(1024, 1) k_smem_random(unsigned* out, int iters) {
  extern __shared__ unsigned s[];
  for (int i = threadIdx.x; i < 48 * 1024; i += blockDim.x) s[i] = (i * 2654435761u) >> 7;
  __syncthreads();
  unsigned a0 = threadIdx.x * 97u, a1 = threadIdx.x * 193u, a2 = threadIdx.x * 389u, a3 = threadIdx.x * 769u;
  for (int it = 0; it < iters; ++it) {
    a0 = s[a0 % (48 * 1024)]; a1 = s[a1 % (48 * 1024)]; a2 = s[a2 % (48 * 1024)]; a3 = s[a3 % (48 * 1024)];
  }
  if ((a0 ^ a1 ^ a2 ^ a3) == 0xdeadbeef) out[0] = a0;
}

// L2-resident streaming read (buffer << L2, re-read many times), 16 B per lane per load
__global__ void k_l2(const uint4* __restrict__ buf, size_t n, unsigned* out, int reps) {
  unsigned acc = 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; ++r)
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      uint4 v;
      asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));
      acc += v.x ^ v.y ^ v.z ^ v.w;
    }
  if (acc == 0xdeadbeef) out[0] = acc;
}

// L2-resident random 32-byte sector gathers (8 B per lane, every lane its own sector)
__global__ void k_l2_gather(const uint2* __restrict__ buf, size_t n_mask, unsigned* out, int iters) {
  unsigned a = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u, acc = 0;
  for (int it = 0; it < iters; ++it) {
    uint2 v;
    asm volatile("ld.global.cg.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(buf + ((a >> 3) & n_mask) * 4));
    acc += v.x;
    a = a * 1664525u + 1013904223u + (v.y & 0);
  }
  if (acc == 0xdeadbeef) out[0] = acc;
}

template <typename F>
static float time_ms(F f, int reps = 5) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  f();
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int i = 0; i < reps; ++i) {
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  return best;
}

}  // namespace

extern "C" int tahoe_measure_peaks(tahoe_peaks_t* out_peaks) {
  using namespace tahoe;
  if (!out_peaks) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  int dev = 0;
  TAHOE_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  TAHOE_CUDA(cudaGetDeviceProperties(&prop, dev));
  const int sms = prop.multiProcessorCount;
  unsigned* out = nullptr;
  uint4* buf = nullptr;
  TAHOE_CUDA(cudaMalloc(&out, 64));
  const size_t smem = 48 * 1024 * 4;
  const size_t bytes = 32u << 20;   // 32 MB: L2-resident
  if (cudaMalloc(&buf, bytes) != cudaSuccess) { cudaFree(out); return cuda_fail(cudaGetLastError(), "cudaMalloc"); }
  cudaMemset(buf, 1, bytes);
  cudaFuncSetAttribute(k_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(k_smem_random, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int iters = 5000;
  float ms = time_ms([&] { k_smem<<<sms, 1024, smem>>>(out, iters); });
  out_peaks->smem_read_gbs = (double)sms * 1024 * iters * 32 * 4 / (ms * 1e-3) / 1e9;
  ms = time_ms([&] { k_smem_random<<<sms, 1024, smem>>>(out, iters * 4); });
  out_peaks->smem_random_word_loads_per_s = (double)sms * 1024 * iters * 4 * 4 / (ms * 1e-3);
  const int reps = 20;
  ms = time_ms([&] { k_l2<<<sms * 8, 512>>>(buf, bytes / 16, out, reps); });
  out_peaks->l2_read_gbs = (double)bytes * reps / (ms * 1e-3) / 1e9;
  const int giters = 2000;
  ms = time_ms([&] { k_l2_gather<<<sms * 8, 256>>>((const uint2*)buf, (bytes / 32) - 1, out, giters); });
  out_peaks->l2_sector_gathers_per_s = (double)sms * 8 * 256 * giters / (ms * 1e-3);
  out_peaks->sm_count = sms;
  out_peaks->l2_bytes = prop.l2CacheSize;
  out_peaks->persisting_l2_max_bytes = prop.persistingL2CacheMaxSize;
  cudaFree(buf); cudaFree(out);
  TAHOE_CUDA(cudaGetLastError());
  return TAHOE_OK;
}
