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
  int depth;
  float missing;
  const uint2* inter;        // [N][tpad]
  const int32_t* orig_ids;   // [N][tpad]
  const int32_t* tree_order; // [tpad]
  float* preds;
  int32_t* leaf_ids;
  Transform tf;
  int warps_per_block;
};

constexpr int kSdGroup = 4;   // tree slots walked together by one lane (ILP over the L2 latency)

// One row, all of this lane's tree slots, kSdGroup at a time.  Returns the lane's partial sum
// (double: lane = tree changes the summation order, so the sum is kept within half an ulp of exact).
__device__ __forceinline__ double sd_walk_row(const SharedDataParams& p, const float* my, size_t row, int lane) {
  double acc = 0.0;
  for (int s0 = lane; s0 < p.tpad; s0 += 32 * kSdGroup) {
    uint32_t n[kSdGroup];
    uint2 e[kSdGroup];
#pragma unroll
    for (int g = 0; g < kSdGroup; ++g) {
      const int slot = s0 + 32 * g;
      n[g] = 0;
      e[g] = slot < p.tpad ? __ldg(p.inter + slot) : make_uint2(0u, kMetaLeaf);
    }
    for (int l = 0; l < p.depth; ++l) {      // a leaf stays where it is; depth bounds every path
      bool any = false;
#pragma unroll
      for (int g = 0; g < kSdGroup; ++g) {
        if (e[g].y & kMetaLeaf) continue;
        const float v = my[e[g].y & kMetaFidMask];
        bool c = branch_right(v, __uint_as_float(e[g].x), e[g].y & kMetaDefLeft, p.missing);
        if (e[g].y & kMetaExchange) c = !c;   // Struct.h:980
        n[g] = 2 * n[g] + 1 + (c ? 1u : 0u);
        e[g] = __ldg(p.inter + (size_t)n[g] * p.tpad + (s0 + 32 * g));
        any = true;
      }
      if (!__any_sync(0xffffffffu, any)) break;
    }
#pragma unroll
    for (int g = 0; g < kSdGroup; ++g) {
      const int slot = s0 + 32 * g;
      if (slot >= p.tpad) continue;
      acc += (double)__uint_as_float(e[g].x);   // padding slots are leaves holding 0.0f
      if (p.leaf_ids) {
        const int t = p.tree_order[slot];
        if (t >= 0) p.leaf_ids[row * (size_t)p.num_trees + t] = p.orig_ids[(size_t)n[g] * p.tpad + slot];
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  return acc;
}

// TMA = true : each warp owns two row buffers; rows arrive by one bulk copy each (cp.async.bulk,
//              row bytes a multiple of 16) completing on the buffer's mbarrier, so the next row
//              streams in from HBM while the current one is traversed.
// TMA = false: the warp copies its row with coalesced loads (any width / alignment).
template <bool TMA>
__global__ void k_shared_data(SharedDataParams p) {
  extern __shared__ __align__(128) uint8_t sd_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const size_t warp_global = (size_t)blockIdx.x * p.warps_per_block + warp;
  const size_t stride = (size_t)gridDim.x * p.warps_per_block;
  if (TMA) {
    const uint32_t row_bytes = (uint32_t)p.cols * 4u;
    const uint32_t s_base = smem_u32(sd_smem);
    const uint32_t bar0 = s_base + (uint32_t)p.warps_per_block * 2u * row_bytes + warp * 16u;
    float* buf0 = reinterpret_cast<float*>(sd_smem + (size_t)warp * 2 * row_bytes);
    if (lane == 0) { mbar_init(bar0, 1); mbar_init(bar0 + 8, 1); fence_mbar_init(); }
    __syncwarp();
    auto load_row = [&](size_t row, int b) {
      if (lane == 0) {
        fence_proxy_async();
        mbar_expect_tx(bar0 + b * 8, row_bytes);
        const uint8_t* src = reinterpret_cast<const uint8_t*>(p.data + row * (size_t)p.cols);
        uint32_t done = 0;
        while (done < row_bytes) {             // pieces of <= 64 KB
          const uint32_t piece = min(row_bytes - done, 65536u);
          bulk_g2s(smem_u32(buf0) + b * row_bytes + done, src + done, piece, bar0 + b * 8);
          done += piece;
        }
      }
    };
    uint32_t phase = 0;
    int b = 0;
    size_t row = warp_global;
    if (row < p.rows) load_row(row, 0);
    for (; row < p.rows; row += stride) {
      __syncwarp();                            // everyone finished reading the buffer being refilled
      if (row + stride < p.rows) load_row(row + stride, b ^ 1);
      mbar_wait(bar0 + b * 8, (phase >> b) & 1u);
      phase ^= 1u << b;
      const float* my = buf0 + (size_t)b * p.cols;
      const double acc = sd_walk_row(p, my, row, lane);
      if (lane == 0) p.preds[row] = apply_transform((float)acc, p.tf);
      b ^= 1;
    }
  } else {
    float* my = reinterpret_cast<float*>(sd_smem) + (size_t)warp * p.cols;
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
      const double acc = sd_walk_row(p, my, row, lane);
      if (lane == 0) p.preds[row] = apply_transform((float)acc, p.tf);
    }
  }
}

int launch_shared_data(tahoe_forest* f, const PredictArgs& a, int* launches) {
  if (f->sparse || !f->d_inter) {
    set_error("SHARED_DATA needs the dense interleaved layout");
    return TAHOE_ERR_NOT_APPLICABLE;
  }
  if (a.rows == 0) return TAHOE_OK;
  const size_t row_bytes = (size_t)f->params.num_cols * sizeof(float);
  SharedDataParams p{};
  p.data = a.data; p.rows = a.rows; p.cols = f->params.num_cols; p.num_trees = f->params.num_trees;
  p.tpad = f->host.tpad; p.depth = f->depth; p.missing = f->params.missing; p.inter = f->d_inter;
  p.orig_ids = f->d_orig_ids; p.tree_order = f->d_tree_order; p.preds = a.preds;
  p.leaf_ids = a.leaf_ids;
  p.tf = Transform{f->params.output, (float)f->params.num_trees, f->params.threshold,
                   f->params.global_bias};
  // bulk copies need 16-byte rows at 16-byte addresses, and two buffers per warp must fit
  const bool tma = (row_bytes % 16 == 0) && ((reinterpret_cast<uintptr_t>(a.data) & 15) == 0) &&
                   row_bytes >= 1024 && 2 * row_bytes + 64 <= (size_t)f->smem_optin;
  auto go = [&](auto kern, int wpb, size_t smem) -> int {
    if (smem > 48 * 1024)
      TAHOE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    TAHOE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, wpb * 32, smem));
    per_sm = std::max(per_sm, 1);
    const size_t want = (a.rows + wpb - 1) / wpb;
    const unsigned grid = (unsigned)std::min<size_t>(want, (size_t)f->sm_count * per_sm);
    p.warps_per_block = wpb;
    kern<<<grid, wpb * 32, smem, a.stream>>>(p);
    TAHOE_CUDA(cudaGetLastError());
    ++*launches;
    return TAHOE_OK;
  };
  if (tma) {
    const int wpb = (int)std::min<size_t>(16, ((size_t)f->smem_optin - 512) / (2 * row_bytes));
    return go(k_shared_data<true>, wpb, (size_t)wpb * 2 * row_bytes + (size_t)wpb * 16 + 64);
  }
  const int wpb = (int)std::min<size_t>(8, (size_t)f->smem_optin / std::max<size_t>(row_bytes, 1));
  if (wpb < 1) {
    set_error("SHARED_DATA: one row does not fit shared memory");
    return TAHOE_ERR_NOT_APPLICABLE;
  }
  return go(k_shared_data<false>, wpb, row_bytes * wpb);
}

}  // namespace tahoe
