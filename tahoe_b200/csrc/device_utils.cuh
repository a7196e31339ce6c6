// device_utils.cuh -- shared device helpers: branch predicate, output transform,
// mbarrier + bulk-copy (TMA, UBLKCP in SASS) wrappers, warp reductions.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace tahoe {

// The reference's branch test (BaseTahoeTest.h:450-452; device Struct.h:976-978):
// float subtract, fabs, compare with float(1e-6); ties go right; NaN goes left.
// Extension (importers): a NaN sentinel means "NaN IS the missing value" -- the convention of the
// trainers (XGBoost, LightGBM, scikit-learn).  In the reference a NaN sentinel matches nothing
// (fabs(v - NaN) <= eps is false) and NaN features fall left; with a finite sentinel this is the
// reference's test unchanged.
__device__ __forceinline__ bool is_missing(float val, float missing) {
  const float eps = 1.0e-6f;
  return (fabsf(val - missing) <= eps) || (missing != missing && val != val);
}
__device__ __forceinline__ bool branch_right(float val, float thr, bool def_left, float missing) {
  return is_missing(val, missing) ? !def_left : (val >= thr);
}

struct Transform {   // forest::predict epilogue, Struct.h:196-209 / BaseTahoeTest.h:467-472
  int output;        // TAHOE_OUT_* flags
  float num_trees;   // the CPU path DIVIDES by num_trees (BaseTahoeTest.h:467); so do we
  float threshold;
  float global_bias;
};

__device__ __forceinline__ float apply_transform(float x, const Transform& t) {
  if (t.output & 0x1) x = x / t.num_trees;
  x += t.global_bias;
  if (t.output & 0x10) x = 1.0f / (1.0f + expf(-x));
  if (t.output & 0x100) x = x > t.threshold ? 1.0f : 0.0f;
  return x;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- shared-space addressing ----------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

// ---- mbarrier + 1-D bulk async copy (global -> shared), the TMA path ---------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}

}  // namespace tahoe
