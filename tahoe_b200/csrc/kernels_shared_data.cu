// kernels_shared_data.cu -- SHARED_DATA (reference K3, Struct.h:992-1191): the sample row
// lives in shared memory and the TREES are spread over the lanes.
//
// The reference gives a whole 256-thread block to one row and finishes with a
// cub::BlockReduce (Struct.h:1015-1017).  Here a WARP owns a row: lane = tree slot in the
// node-major interleaved layout [node][tree] (adjacent lanes read adjacent words while
// they sit on the same node -- which the probability reorder and the simhash tree order
// make the common case), per-lane partial sums over the tree groups, one shuffle
// reduction per row.  No block barrier in the row loop, no idle lanes when T < 256.
// This is the strategy for wide rows (SVHN-like: a 12 KB row is read once, coalesced)
// and for small batches, where trees are the only parallelism.
#include "device_utils.cuh"
#include "forest.h"

namespace tahoe {

struct SharedDataParams {
  const float* data;
  size_t rows;
  int cols;
  int num_trees;
  int tpad;
  float missing;
  const uint2* inter;        // [N][tpad]
  const int32_t* orig_ids;   // [N][tpad]
  const int32_t* tree_order; // [tpad]
  float* preds;
  int32_t* leaf_ids;
  Transform tf;
  int warps_per_block;
};

__global__ void k_shared_data(SharedDataParams p) {
  extern __shared__ float s_row[];  // [warps_per_block][cols]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* my = s_row + (size_t)warp * p.cols;
  const size_t warp_global = (size_t)blockIdx.x * p.warps_per_block + warp;
  const size_t stride = (size_t)gridDim.x * p.warps_per_block;
  const bool vec4 = ((p.cols & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.data) & 15) == 0);
  for (size_t row = warp_global; row < p.rows; row += stride) {
    const float* g = p.data + row * (size_t)p.cols;
    __syncwarp();
    if (vec4) {
      const float4* g4 = reinterpret_cast<const float4*>(g);
      float4* m4 = reinterpret_cast<float4*>(my);
      for (int i = lane; i < (p.cols >> 2); i += 32) m4[i] = __ldg(g4 + i);
    } else {
      for (int i = lane; i < p.cols; i += 32) my[i] = __ldg(g + i);
    }
    __syncwarp();
    double acc = 0.0;   // lane = tree changes the summation order; double keeps the result within half an ulp of exact
    for (int slot = lane; slot < p.tpad; slot += 32) {
      uint32_t n = 0;
      uint2 e = __ldg(p.inter + slot);
      while (!(e.y & kMetaLeaf)) {
        const float v = my[e.y & kMetaFidMask];
        bool c = branch_right(v, __uint_as_float(e.x), e.y & kMetaDefLeft, p.missing);
        if (e.y & kMetaExchange) c = !c;      // Struct.h:980
        n = 2 * n + 1 + (c ? 1u : 0u);
        e = __ldg(p.inter + (size_t)n * p.tpad + slot);
      }
      acc += (double)__uint_as_float(e.x);    // padding slots are leaves holding 0.0f
      if (p.leaf_ids) {
        const int t = p.tree_order[slot];
        if (t >= 0) p.leaf_ids[row * (size_t)p.num_trees + t] = p.orig_ids[(size_t)n * p.tpad + slot];
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) p.preds[row] = apply_transform((float)acc, p.tf);
  }
}

int launch_shared_data(tahoe_forest* f, const PredictArgs& a, int* launches) {
  if (f->sparse || !f->d_inter) {
    set_error("SHARED_DATA needs the dense interleaved layout");
    return TAHOE_ERR_NOT_APPLICABLE;
  }
  if (a.rows == 0) return TAHOE_OK;
  const size_t row_bytes = (size_t)f->params.num_cols * sizeof(float);
  int wpb = (int)std::min<size_t>(8, (size_t)f->smem_optin / std::max<size_t>(row_bytes, 1));
  if (wpb < 1) {
    set_error("SHARED_DATA: one row does not fit shared memory");
    return TAHOE_ERR_NOT_APPLICABLE;
  }
  SharedDataParams p{};
  p.data = a.data; p.rows = a.rows; p.cols = f->params.num_cols; p.num_trees = f->params.num_trees;
  p.tpad = f->host.tpad; p.missing = f->params.missing; p.inter = f->d_inter;
  p.orig_ids = f->d_orig_ids; p.tree_order = f->d_tree_order; p.preds = a.preds;
  p.leaf_ids = a.leaf_ids; p.warps_per_block = wpb;
  p.tf = Transform{f->params.output, (float)f->params.num_trees, f->params.threshold,
                   f->params.global_bias};
  const size_t smem = row_bytes * wpb;
  // vectorised staging needs each warp's buffer 16-byte aligned
  if (smem > 48 * 1024)
    TAHOE_CUDA(cudaFuncSetAttribute(k_shared_data, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  TAHOE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_shared_data, wpb * 32, smem));
  per_sm = std::max(per_sm, 1);
  const size_t want = (a.rows + wpb - 1) / wpb;
  const unsigned grid = (unsigned)std::min<size_t>(want, (size_t)f->sm_count * per_sm);
  k_shared_data<<<grid, wpb * 32, smem, a.stream>>>(p);
  TAHOE_CUDA(cudaGetLastError());
  ++*launches;
  return TAHOE_OK;
}

}  // namespace tahoe
