// gen.cu -- seeded synthetic sample matrix generated directly in device memory.
// Bit-identical to tahoe_b200/synth.py gen_data (same splitmix64 counter RNG), so a
// test can check any slice of a 5M-row device matrix against the oracle run on the
// same rows generated on the host.  Replaces generate_data (BaseTahoeTest.h:404-438),
// whose rand()-based stream is not reproducible across libcs; U(-1,1) law kept (:123-130).
#include "forest.h"

namespace tahoe {

__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__global__ void k_gen_data(uint64_t key, size_t idx0, size_t n, uint32_t miss_cut, float missing,
                           float* out) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t r = splitmix64((uint64_t)(idx0 + i) ^ key);
    float v = (float)(uint32_t)(r >> 40) * (1.0f / 8388608.0f) - 1.0f;
    if ((uint32_t)(r & 0xFFFFFFull) < miss_cut) v = missing;
    out[i] = v;
  }
}

}  // namespace tahoe

extern "C" int tahoe_gen_data_device(uint64_t seed, size_t row0, size_t num_rows, int num_cols,
                                     float miss_frac, float missing, float* d_out, void* stream) {
  using namespace tahoe;
  if (!d_out || num_cols < 1) { set_error("tahoe_gen_data_device: bad arguments"); return TAHOE_ERR_INVALID; }
  const size_t n = num_rows * (size_t)num_cols;
  if (n == 0) return TAHOE_OK;
  const int kDataStream = 1;  // synth.S_DATA
  const uint64_t key = splitmix64(seed + 0x632BE59BD9B4E019ull * (uint64_t)(kDataStream + 1));
  const uint32_t miss_cut = miss_frac > 0.0f ? (uint32_t)(int)((double)miss_frac * 16777216.0) : 0u;
  const unsigned grid = (unsigned)std::min<size_t>((n + 255) / 256, 148 * 32);
  k_gen_data<<<grid, 256, 0, (cudaStream_t)stream>>>(key, row0 * (size_t)num_cols, n, miss_cut, missing, d_out);
  TAHOE_CUDA(cudaGetLastError());
  return TAHOE_OK;
}
