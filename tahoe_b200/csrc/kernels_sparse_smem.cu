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
//                   three independent tree sequences per lane for latency.  A walk that steps past the
//                   cut is parked in a per-warp list; whenever 32 are parked the warp finishes them
//                   together in L2 on the full breadth-first copy, one walk per lane, and hands the
//                   leaf values back to the owning lanes with shuffles.
//
// The sum of a row is therefore NOT in tree order (two sequences per lane + parked walks): scores
// agree with the CPU within the float-reorder tolerance, leaf ids are exact.
#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "device_utils.cuh"
#include "forest.h"

namespace tahoe {

struct SparseSmemParams {
  const float* tiles;          // [tiles][cols][32]
  const uint32_t* tile_flags;  // [tiles] 1 = some feature of the tile equals the missing-value sentinel
  const float* data;           // the caller's rows (features of the last parked walks of a tile)
  size_t num_tiles, rows;
  int cols;
  const uint2* top;            // [chunks * tpc][cut]
  const uint2* full;           // breadth-first S8 nodes
  const int64_t* tree_offsets; // [T+1]
  const int32_t* orig;         // breadth-first index -> caller's node index
  int num_trees, tpc, num_chunks, cut, fid_bits;
  float missing;
  float* preds;                // partial sums between chunk passes
  float* out;                  // final scores (last chunk pass); may be peer-mapped
  int32_t* leaf_ids;
  Transform tf;
};

// Shared-memory image of a tree top (packer: sp_top), 8 bytes per node:
//   x = threshold / leaf value
//   y = (child delta * 8) << 17 | fid << 11 | def_left << 1 | leaf
// so that delta*8 = y >> 17 is the byte step to the left child and y & 0x1F800 is already the byte
// offset of the feature's column row in the tile region (2048 bytes per column).
constexpr uint32_t kTopLeaf = 1u, kTopDefLeft = 2u, kTopColMask = 0x1F800u;

__global__ void __launch_bounds__(256) k_float_tiles(const float* __restrict__ data, size_t rows, int cols,
                                                     float missing, float* __restrict__ tiles,
                                                     uint32_t* __restrict__ tile_flags, size_t num_tiles) {
  const int lane = threadIdx.x & 31;
  const size_t warp = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const size_t stride = (size_t)gridDim.x * (blockDim.x >> 5);
  for (size_t tile = warp; tile < num_tiles; tile += stride) {
    const size_t row = tile * 32 + lane;
    const float* src = data + row * (size_t)cols;
    float* dst = tiles + tile * (size_t)cols * 32 + lane;
    bool miss = false;
    for (int c0 = 0; c0 < cols; c0 += 8) {
      float v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = (row < rows && c0 + k < cols) ? __ldg(src + c0 + k) : 0.0f;
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (c0 + k < cols) {
          dst[(size_t)(c0 + k) * 32] = v[k];
          miss |= is_missing(v[k], missing);
        }
    }
    const bool any = __any_sync(0xffffffffu, miss);
    if (lane == 0) tile_flags[tile] = any ? 1u : 0u;     // tiles without the sentinel take the short compare
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
//   [.., + tpc * cut * 8 + 16)      tops of the chunk's trees, then one sentinel leaf
//   [.., + 16 * kSparsePend * 8)    parked walks per warp: {tree << 5 | lane, node}
//   [.., + 17 * 8)                  mbarriers: chunk, one tile buffer per warp
template <bool LEAF_IDS, int G>
__global__ void __launch_bounds__(kSparseWarps * 32, 1) k_sparse_smem(SparseSmemParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t kColStride = kSparseWarps * 128u;
  static_assert(kColStride == 2048u, "kTopColMask assumes 2048-byte column rows");
  const uint32_t s_base = smem_u32(smem);
  const uint32_t tile_bytes = (uint32_t)p.cols * 128u;
  const uint32_t forest_off = (uint32_t)p.cols * kColStride;
  const uint32_t cut8 = (uint32_t)p.cut * 8u;
  const uint32_t chunk_bytes = (uint32_t)p.tpc * cut8;
  const uint32_t pend_off = forest_off + chunk_bytes + 16u;   // a sentinel leaf follows the last tree
  const uint32_t bars_off = pend_off + kSparseWarps * kSparsePend * 8u;
  const uint32_t bar_chunk = s_base + bars_off;
  const uint32_t bar_tile = bar_chunk + 8 + warp * 8;
  uint2* pend = reinterpret_cast<uint2*>(smem + pend_off) + warp * kSparsePend;
  const uint32_t s_forest = s_base + forest_off;
  const uint32_t slot = s_base + (uint32_t)warp * 128u;       // this warp's row of every column
  const uint32_t slot_lane_rel = (uint32_t)warp * 128u + lane * 4u;   // < 2048: disjoint from the column field
  const uint32_t fid_mask = (1u << p.fid_bits) - 1u;
  const uint32_t lt_mask = (1u << lane) - 1u;

  if (threadIdx.x == 0) {
    mbar_init(bar_chunk, 1);
    for (int i = 0; i < kSparseWarps; ++i) mbar_init(bar_chunk + 8 + i * 8, 1);
    fence_mbar_init();
    // a sequence that has finished its trees keeps reading the node after them: make that a leaf
    // (its feature address must stay inside the tile region)
    *reinterpret_cast<uint2*>(smem + forest_off + chunk_bytes) = make_uint2(0u, kTopLeaf);
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
    // G free-running tree sequences per lane: sequence g walks trees [g*per, min((g+1)*per, ntrees))
    const int per = (ntrees + G - 1) / G;

    for (; tile < p.num_tiles; tile += warp_stride) {
      const size_t row = tile * 32 + lane;
      const bool live = row < p.rows;
      float acc = (!first && live) ? p.preds[row] : 0.0f;
      const bool has_missing = __ldg(p.tile_flags + tile) != 0;
      mbar_wait(bar_tile, tile_phase);
      tile_phase ^= 1;

      int count = 0;                       // parked walks of this warp (warp-uniform)
      float acc_deep = 0.0f;
      // Finish k parked walks (the newest ones), one per lane, in L2 on the full breadth-first copy.
      // FROM_TILE: features come from this warp's tile; else from the caller's rows in global memory
      // (the last drain of a tile, so that the tile buffer can be refilled underneath it).
      auto drain = [&](int k, auto from_tile) {
        constexpr bool FROM_TILE = decltype(from_tile)::value;
        __syncwarp();
        uint2 ent = make_uint2(0u, 0u);
        float val = 0.0f;
        if (lane < k) {
          ent = pend[count - k + lane];
          const uint32_t tree = ent.x >> 5;
          const int64_t off = p.tree_offsets[tree];
          const uint2* base = p.full + off;
          const uint32_t my_row = slot + (ent.x & 31u) * 4u;
          const size_t r = tile * 32 + (ent.x & 31u);          // < rows: only live lanes park
          const float* grow = p.data + r * (size_t)p.cols;
          uint32_t n = ent.y;
          uint2 e = __ldg(base + n);
          while (!(e.y & kMetaLeaf)) {
            const uint32_t fid = e.y & fid_mask;
            const float v = FROM_TILE ? lds_f32(my_row + fid * kColStride) : __ldg(grow + fid);
            n += ((e.y & 0x3FFFFFFFu) >> p.fid_bits) +
                 (branch_right(v, __uint_as_float(e.x), e.y & kMetaDefLeft, p.missing) ? 1u : 0u);
            e = __ldg(base + n);
          }
          val = __uint_as_float(e.x);
          if (LEAF_IDS) p.leaf_ids[r * (size_t)p.num_trees + tree] = p.orig[off + n];
        }
        __syncwarp();
        for (int j = 0; j < k; ++j) {      // parked order: the sum is reproducible
          const uint32_t r = __shfl_sync(0xffffffffu, ent.x, j) & 31u;
          const float v = __shfl_sync(0xffffffffu, val, j);
          if (r == (uint32_t)lane) acc_deep += v;
        }
        count -= k;
      };

      // per sequence: a = shared address of the current node, lim = end of the current tree's top
      // (= base of the next tree), rem = trees left (0 for rows past the end).  The step is written
      // branch-free: selects and predicated adds only, one warp vote for the (rare) parking.
      uint32_t a[G], lim[G];
      int rem[G], tlast[G];                // tlast - rem = index of the current tree in the forest
      float accg[G];
      uint2 e[G];
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const int t0 = g * per;
        rem[g] = live ? max(0, min(per, ntrees - t0)) : 0;
        tlast[g] = tree0 + t0 + rem[g];
        a[g] = s_forest + (uint32_t)min(t0, ntrees - 1) * cut8;
        lim[g] = a[g] + cut8;
        accg[g] = 0.0f;
        e[g] = lds_u64(a[g]);
      }
      auto sequences = [&](auto missing_tag) {
        constexpr bool MISSING = decltype(missing_tag)::value;
        while (true) {
          int left = 0;
#pragma unroll
          for (int g = 0; g < G; ++g) left |= rem[g];          // rem >= 0
          if (!__any_sync(0xffffffffu, left != 0)) break;
          bool deep[G];
          uint32_t a2[G];
          bool any_deep = false;
#pragma unroll
          for (int g = 0; g < G; ++g) {
            const bool act = rem[g] != 0;
            const bool leaf = (e[g].y & kTopLeaf) != 0;
            const float v = lds_f32(s_base + ((e[g].y & kTopColMask) | slot_lane_rel));   // a leaf reads column 0
            bool c = v >= __uint_as_float(e[g].x);
            if (MISSING) {
              if (is_missing(v, p.missing)) c = !(e[g].y & kTopDefLeft);   // BaseTahoeTest.h:451-452
            }
            a2[g] = a[g] + (e[g].y >> 17) + (c ? 8u : 0u);
            const bool past = a2[g] >= lim[g];
            const bool done = act & (leaf | past);
            deep[g] = act & !leaf & past;
            any_deep |= deep[g];
            accg[g] += (act & leaf) ? __uint_as_float(e[g].x) : 0.0f;
            if (LEAF_IDS) {
              if (act & leaf) {
                const int tr = tlast[g] - rem[g];
                p.leaf_ids[row * (size_t)p.num_trees + tr] = p.orig[p.tree_offsets[tr] + ((a[g] - (lim[g] - cut8)) >> 3)];
              }
            }
            // next node of the sequence (the parked walk, if any, is recorded below from a2 / lim / rem)
            const uint32_t an = done ? lim[g] : (act ? a2[g] : a[g]);
            e[g] = lds_u64(an);            // a finished sequence sits on the next tree's root or the sentinel leaf
            a[g] = an;
          }
          // one vote for all sequences; parking is the rare path
          if (__any_sync(0xffffffffu, any_deep)) {
#pragma unroll
            for (int g = 0; g < G; ++g) {
              const uint32_t park = __ballot_sync(0xffffffffu, deep[g]);
              if (deep[g])
                pend[count + __popc(park & lt_mask)] =
                    make_uint2(((uint32_t)(tlast[g] - rem[g]) << 5) | lane, (a2[g] - (lim[g] - cut8)) >> 3);
              count += __popc(park);
            }
          }
#pragma unroll
          for (int g = 0; g < G; ++g) {
            const bool done = a[g] == lim[g];          // the sequence moved on to the next tree
            lim[g] += done ? cut8 : 0u;
            rem[g] -= done ? 1 : 0;
          }
          while (count >= 32) drain(32, std::true_type{});   // keeps count + 32*G within kSparsePend
        }
      };
      if (has_missing) sequences(std::true_type{}); else sequences(std::false_type{});
      // the tile is consumed: refill this warp's buffer underneath the last drain
      __syncwarp();
      const size_t next = tile + warp_stride;
      if (next < p.num_tiles) load_tile(next);
      if (count > 0) drain(count, std::false_type{});
#pragma unroll
      for (int g = 0; g < G; ++g) acc += accg[g];
      acc += acc_deep;
      if (live) {
        if (last) p.out[row] = apply_transform(acc, p.tf);
        else p.preds[row] = acc;
      }
    }
  }
}

int launch_float_tiles(tahoe_forest* f, const float* data, size_t rows, float* tiles_out, uint32_t* tile_flags, size_t tiles,
                       cudaStream_t stream) {
  const unsigned grid = (unsigned)std::min<size_t>((tiles + 7) / 8, (size_t)f->sm_count * 8);
  k_float_tiles<<<grid, 256, 0, stream>>>(data, rows, f->params.num_cols, f->params.missing, tiles_out, tile_flags, tiles);
  TAHOE_CUDA(cudaGetLastError());
  return TAHOE_OK;
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
    k_float_tiles<<<grid, 256, 0, a.stream>>>(a.data, a.rows, f->params.num_cols, f->params.missing, ftiles,
                                              a.ws->tile_flags, tiles);
    TAHOE_CUDA(cudaGetLastError());
    ++*launches;
  }
  if (ev) TAHOE_CUDA(cudaEventRecord(ev[1], a.stream));
  SparseSmemParams p{};
  p.tiles = ftiles; p.tile_flags = a.ws->tile_flags; p.data = a.data; p.num_tiles = tiles; p.rows = a.rows; p.cols = f->params.num_cols;
  p.top = f->d_sp_top; p.full = f->d_sp_full; p.tree_offsets = f->d_tree_offsets; p.orig = f->d_sp_orig;
  p.num_trees = f->params.num_trees; p.tpc = f->sp_tpc; p.num_chunks = f->sp_chunks; p.cut = f->sp_cut;
  p.fid_bits = f->host.sparse_fid_bits; p.missing = f->params.missing;
  p.preds = a.preds; p.out = a.out; p.leaf_ids = a.leaf_ids;
  p.tf = Transform{f->params.output, (float)f->params.num_trees, f->params.threshold, f->params.global_bias};
  unsigned grid = (unsigned)std::min<size_t>((size_t)f->sm_count, (tiles + kSparseWarps - 1) / kSparseWarps);
  if (f->tune.sparse_grid > 0) grid = std::max(1u, std::min(grid, (unsigned)f->tune.sparse_grid));
  auto go = [&](auto kern) -> int {
    TAHOE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f->sp_smem));
    kern<<<grid, kSparseWarps * 32, f->sp_smem, a.stream>>>(p);
    TAHOE_CUDA(cudaGetLastError());
    ++*launches;
    return TAHOE_OK;
  };
  const int group = f->tune.sparse_group;   // default 3: measured on config 4: 28.7 / 24.6 / 23.4 / 24.5 ms for 1 / 2 / 3 / 4 sequences per lane
  if (a.leaf_ids) rc = go(k_sparse_smem<true, 2>);
  else if (group == 4) rc = go(k_sparse_smem<false, 4>);
  else rc = group == 3 ? go(k_sparse_smem<false, 3>) : (group == 1 ? go(k_sparse_smem<false, 1>) : go(k_sparse_smem<false, 2>));
  if (ev && rc == TAHOE_OK) TAHOE_CUDA(cudaEventRecord(ev[2], a.stream));
  return rc;
}

}  // namespace tahoe
