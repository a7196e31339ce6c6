// kernels_sparse_smem.cu -- SHARED_FOREST / SPLIT_SHARED_FOREST for SPARSE (compact, unbalanced)
// forests: BASELINE config 4, "deep unbalanced forest exceeding shared memory".  The reference has
// no such path (its sparse forest is disabled, BaseTahoeTest.h:727-846; its shared-forest kernels
// are dense only, Struct.h:1245-1606).
//
// Deep trees do not fit shared memory, but walks are short: most leave the tree within a few levels
// of the root.  So every tree is renumbered breadth-first (packer: T8) and a chunk keeps only the
// first `cut` nodes of each of its trees in shared memory:
//
//   k_float_tiles   rows -> column-major 32-row tiles [tile][col][32 lanes] (lane i only ever touches
//                   bank i), once per predict
//   k_sparse_smem   persistent CTAs, one per SM.  Per chunk: bulk-copy the tree tops, then stream
//                   every row tile through them.  thread = sample.  Lanes are FREE-RUNNING: a lane
//                   that reaches a leaf starts its next tree at once instead of idling until the
//                   warp's deepest walk ends (in shared memory there is no sector sharing to lose),
//                   two independent tree sequences per lane for latency.  A walk that steps past the
//                   cut is parked in a per-warp list; whenever 32 are parked the warp finishes them
//                   together in L2 on the full breadth-first copy, one walk per lane, and hands the
//                   leaf values back to the owning lanes with shuffles.
//
// The sum of a row is therefore NOT in tree order (two sequences per lane + parked walks): scores
// agree with the CPU within the float-reorder tolerance, leaf ids are exact.
#include <algorithm>
#include <cstdlib>

#include "device_utils.cuh"
#include "forest.h"

namespace tahoe {

struct SparseSmemParams {
  const float* tiles;          // [tiles][cols][32]
  size_t num_tiles, rows;
  int cols;
  const uint2* top;            // [chunks * tpc][cut]
  const uint2* full;           // breadth-first S8 nodes
  const int64_t* tree_offsets; // [T+1]
  const int32_t* orig;         // breadth-first index -> caller's node index
  int num_trees, tpc, num_chunks, cut, fid_bits;
  float missing;
  float* preds;
  int32_t* leaf_ids;
  Transform tf;
};

__global__ void __launch_bounds__(256) k_float_tiles(const float* __restrict__ data, size_t rows, int cols,
                                                     float* __restrict__ tiles, size_t num_tiles) {
  const int lane = threadIdx.x & 31;
  const size_t warp = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const size_t stride = (size_t)gridDim.x * (blockDim.x >> 5);
  for (size_t tile = warp; tile < num_tiles; tile += stride) {
    const size_t row = tile * 32 + lane;
    const float* src = data + row * (size_t)cols;
    float* dst = tiles + tile * (size_t)cols * 32 + lane;
    for (int c0 = 0; c0 < cols; c0 += 8) {
      float v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = (row < rows && c0 + k < cols) ? __ldg(src + c0 + k) : 0.0f;
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (c0 + k < cols) dst[(size_t)(c0 + k) * 32] = v[k];
    }
  }
}

__device__ __forceinline__ uint2 lds_u64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}

// Shared-memory map (byte offsets from smem[]):
//   [0, cols * 2048)                tile region: word (col, warp, lane) at col*2048 + warp*128 + lane*4
//   [.., + tpc * cut * 8)           tops of the chunk's trees
//   [.., + 16 * kSparsePend * 8)    parked walks per warp: {tree << 5 | lane, node}
//   [.., + 17 * 8)                  mbarriers: chunk, one tile buffer per warp
template <bool LEAF_IDS>
__global__ void __launch_bounds__(kSparseWarps * 32, 1) k_sparse_smem(SparseSmemParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t kColStride = kSparseWarps * 128u;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t tile_bytes = (uint32_t)p.cols * 128u;
  const uint32_t forest_off = (uint32_t)p.cols * kColStride;
  const uint32_t chunk_bytes = (uint32_t)p.tpc * (uint32_t)p.cut * 8u;
  const uint32_t pend_off = forest_off + chunk_bytes;
  const uint32_t bars_off = pend_off + kSparseWarps * kSparsePend * 8u;
  const uint32_t bar_chunk = s_base + bars_off;
  const uint32_t bar_tile = bar_chunk + 8 + warp * 8;
  uint2* pend = reinterpret_cast<uint2*>(smem + pend_off) + warp * kSparsePend;
  const uint32_t s_forest = s_base + forest_off;
  const uint32_t slot = s_base + (uint32_t)warp * 128u;       // this warp's row of every column
  const uint32_t slot_lane = slot + lane * 4u;
  const uint32_t fid_mask = (1u << p.fid_bits) - 1u;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t ucut = (uint32_t)p.cut;

  if (threadIdx.x == 0) {
    mbar_init(bar_chunk, 1);
    for (int i = 0; i < kSparseWarps; ++i) mbar_init(bar_chunk + 8 + i * 8, 1);
    fence_mbar_init();
  }
  __syncthreads();

  const size_t warp_global = (size_t)blockIdx.x * kSparseWarps + warp;
  const size_t warp_stride = (size_t)gridDim.x * kSparseWarps;
  uint32_t tile_phase = 0, chunk_phase = 0;

  auto load_tile = [&](size_t tile) {      // one 128-byte bulk copy per column row
    if (lane == 0) mbar_expect_tx(bar_tile, tile_bytes);
    __syncwarp();
    const float* src = p.tiles + tile * (size_t)p.cols * 32;
    for (int c = lane; c < p.cols; c += 32) {
      fence_proxy_async();
      bulk_g2s(slot + c * kColStride, src + c * 32, 128u, bar_tile);
    }
  };

  for (int chunk = 0; chunk < p.num_chunks; ++chunk) {
    const int tree0 = chunk * p.tpc;
    const int ntrees = min(p.tpc, p.num_trees - tree0);
    __syncthreads();                       // everyone is done with the previous chunk's trees
    if (threadIdx.x == 0) {
      fence_proxy_async();
      mbar_expect_tx(bar_chunk, chunk_bytes);
      const uint8_t* src = reinterpret_cast<const uint8_t*>(p.top + (size_t)tree0 * p.cut);
      uint32_t done = 0;
      while (done < chunk_bytes) {
        const uint32_t piece = min(chunk_bytes - done, 65536u);
        bulk_g2s(s_forest + done, src + done, piece, bar_chunk);
        done += piece;
      }
    }
    size_t tile = warp_global;
    if (tile < p.num_tiles) load_tile(tile);
    mbar_wait(bar_chunk, chunk_phase);
    chunk_phase ^= 1;
    const bool first = chunk == 0, last = chunk == p.num_chunks - 1;

    for (; tile < p.num_tiles; tile += warp_stride) {
      const size_t row = tile * 32 + lane;
      const bool live = row < p.rows;
      float acc = (!first && live) ? p.preds[row] : 0.0f;
      mbar_wait(bar_tile, tile_phase);
      tile_phase ^= 1;

      int count = 0;                       // parked walks of this warp (warp-uniform)
      float acc_deep = 0.0f;
      // finish k parked walks (the newest ones), one per lane, in L2
      auto drain = [&](int k) {
        __syncwarp();
        uint2 ent = make_uint2(0u, 0u);
        float val = 0.0f;
        if (lane < k) {
          ent = pend[count - k + lane];
          const uint32_t tree = ent.x >> 5;
          const int64_t off = p.tree_offsets[tree];
          const uint2* base = p.full + off;
          const uint32_t my_row = slot + (ent.x & 31u) * 4u;
          uint32_t n = ent.y;
          uint2 e = __ldg(base + n);
          while (!(e.y & kMetaLeaf)) {
            const float v = lds_f32(my_row + (e.y & fid_mask) * kColStride);
            n += ((e.y & 0x3FFFFFFFu) >> p.fid_bits) +
                 (branch_right(v, __uint_as_float(e.x), e.y & kMetaDefLeft, p.missing) ? 1u : 0u);
            e = __ldg(base + n);
          }
          val = __uint_as_float(e.x);
          if (LEAF_IDS) {
            const size_t r = tile * 32 + (ent.x & 31u);
            if (r < p.rows) p.leaf_ids[r * (size_t)p.num_trees + tree] = p.orig[off + n];
          }
        }
        __syncwarp();
        for (int j = 0; j < k; ++j) {      // parked order: the sum is reproducible
          const uint32_t r = __shfl_sync(0xffffffffu, ent.x, j) & 31u;
          const float v = __shfl_sync(0xffffffffu, val, j);
          if (r == (uint32_t)lane) acc_deep += v;
        }
        count -= k;
      };

      // two free-running tree sequences per lane: [0, half) and [half, ntrees)
      const int half = (ntrees + 1) >> 1;
      int t[2] = {0, half};
      const int tend[2] = {half, ntrees};
      uint32_t n[2] = {0u, 0u};
      uint2 e[2];
      float accg[2] = {0.0f, 0.0f};
#pragma unroll
      for (int g = 0; g < 2; ++g) e[g] = lds_u64(s_forest + (uint32_t)min(t[g], ntrees - 1) * ucut * 8u);
      while (__any_sync(0xffffffffu, t[0] < tend[0] || t[1] < tend[1])) {
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const bool act = t[g] < tend[g];
          const bool leaf = (e[g].y & kMetaLeaf) != 0;
          const float v = lds_f32(slot_lane + (e[g].y & fid_mask) * kColStride);   // a leaf reads column 0
          const uint32_t nn = n[g] + ((e[g].y & 0x3FFFFFFFu) >> p.fid_bits) +
                              (branch_right(v, __uint_as_float(e[g].x), e[g].y & kMetaDefLeft, p.missing) ? 1u : 0u);
          const bool deep = !leaf && nn >= ucut;
          const bool done = leaf || deep;
          if (act && leaf) {
            accg[g] += __uint_as_float(e[g].x);
            if (LEAF_IDS && live)
              p.leaf_ids[row * (size_t)p.num_trees + tree0 + t[g]] = p.orig[p.tree_offsets[tree0 + t[g]] + n[g]];
          }
          const uint32_t park = __ballot_sync(0xffffffffu, act && deep);
          if (park) {
            if (act && deep) pend[count + __popc(park & lt_mask)] = make_uint2(((uint32_t)(tree0 + t[g]) << 5) | lane, nn);
            count += __popc(park);
          }
          if (act) {
            t[g] += done ? 1 : 0;
            n[g] = done ? 0u : nn;
          }
          const int tt = min(t[g], tend[g] - 1);       // a finished sequence keeps re-reading a valid node
          e[g] = lds_u64(s_forest + ((uint32_t)max(tt, 0) * ucut + n[g]) * 8u);
        }
        if (count >= 32) drain(32);
      }
      if (count > 0) drain(count);
      // the tile is consumed: refill this warp's buffer, then finish the row
      __syncwarp();
      const size_t next = tile + warp_stride;
      if (next < p.num_tiles) load_tile(next);
      acc = acc + accg[0] + accg[1] + acc_deep;
      if (live) p.preds[row] = last ? apply_transform(acc, p.tf) : acc;
    }
  }
}

int launch_sparse_smem(tahoe_forest* f, const PredictArgs& a, int* launches) {
  if (!f->sparse || f->sp_tpc < 1 || !f->d_sp_top) {
    set_error("shared-memory strategies on a sparse forest need 8-byte breadth-first nodes and room for "
              "one tile per warp next to the tree tops");
    return TAHOE_ERR_NOT_APPLICABLE;
  }
  if (a.rows == 0) return TAHOE_OK;
  int rc = ensure_workspace(f, a.ws, a.rows);
  if (rc) return rc;
  const size_t tiles = (a.rows + 31) / 32;
  float* ftiles = reinterpret_cast<float*>(a.ws->qtiles);
  cudaEvent_t* ev = nullptr;
  if (f->profile && a.ws == &f->ws[0] && f->prof_count < kProfRing) {
    ev = &f->prof_events[(size_t)f->prof_count * 3];
    ++f->prof_count;
    TAHOE_CUDA(cudaEventRecord(ev[0], a.stream));
  }
  {
    const unsigned grid = (unsigned)std::min<size_t>((tiles + 7) / 8, (size_t)f->sm_count * 8);
    k_float_tiles<<<grid, 256, 0, a.stream>>>(a.data, a.rows, f->params.num_cols, ftiles, tiles);
    TAHOE_CUDA(cudaGetLastError());
    ++*launches;
  }
  if (ev) TAHOE_CUDA(cudaEventRecord(ev[1], a.stream));
  SparseSmemParams p{};
  p.tiles = ftiles; p.num_tiles = tiles; p.rows = a.rows; p.cols = f->params.num_cols;
  p.top = f->d_sp_top; p.full = f->d_sp_full; p.tree_offsets = f->d_tree_offsets; p.orig = f->d_sp_orig;
  p.num_trees = f->params.num_trees; p.tpc = f->sp_tpc; p.num_chunks = f->sp_chunks; p.cut = f->sp_cut;
  p.fid_bits = f->host.sparse_fid_bits; p.missing = f->params.missing;
  p.preds = a.preds; p.leaf_ids = a.leaf_ids;
  p.tf = Transform{f->params.output, (float)f->params.num_trees, f->params.threshold, f->params.global_bias};
  const unsigned grid = (unsigned)std::min<size_t>((size_t)f->sm_count, (tiles + kSparseWarps - 1) / kSparseWarps);
  auto go = [&](auto kern) -> int {
    TAHOE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f->sp_smem));
    kern<<<grid, kSparseWarps * 32, f->sp_smem, a.stream>>>(p);
    TAHOE_CUDA(cudaGetLastError());
    ++*launches;
    return TAHOE_OK;
  };
  rc = a.leaf_ids ? go(k_sparse_smem<true>) : go(k_sparse_smem<false>);
  if (ev && rc == TAHOE_OK) TAHOE_CUDA(cudaEventRecord(ev[2], a.stream));
  return rc;
}

}  // namespace tahoe
