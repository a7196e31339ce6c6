// kernels_smem_traverse.cuh -- k_smem_forest: SHARED_FOREST / SPLIT_SHARED_FOREST traversal
// (reference K4/K5, Struct.h:1245-1409 and 1414-1606).  Included by kernels_smem_depth.cu, which is
// compiled once per tree depth (-DTAHOE_DEPTH=N) so that the 14 x 28 instantiations build in parallel.
#pragma once
#include "device_utils.cuh"
#include "forest.h"

namespace tahoe {

// ---------------------------------------------------------------------------------------------
struct SmemParams {
  const uint32_t* qtiles;
  const uint32_t* tile_flags;
  size_t num_tiles;
  size_t rows;
  int cols;
  const uint32_t* qwords;   // [padded_trees][2^(d+1)]
  int num_trees;            // real trees (leaf ids are written for these only)
  int padded_trees;         // trees in qwords: whole traversal groups, the extra ones all zero
  int trees_per_chunk, num_chunks;
  int tree_base;            // index (in the whole padded forest) of this launch's first tree: qwords points at it
  int first_pass, last_pass;  // this launch holds the forest's first / last chunk (partial sums start / end here)
  int top_levels;           // levels of every tree decided on the TopWords kernel parameter (0 or 2): picks the instantiation
  int cluster;              // > 1: k_smem_forest_cluster with this many CTAs per cluster (kernels_smem_cluster.cuh)
  uint32_t col_mask, miss_code;
  // tile region: word (col, slot, lane) at col*col_stride + slot*slot_stride + lane*4, and
  // (node & col_mask) == col*col_stride, so a lane's address is one AND-OR.  Two layouts:
  //   wide    column-major  col_stride = 128*slots, slot_stride = 128   (column field at bit 7+log2 slots)
  //   compact slot-major    col_stride = 128, slot_stride = 128 * cols rounded up to a power of two
  //           (column field at bit 7: for forests whose rank + column bits leave no room for the wide field)
  uint32_t col_stride, slot_stride;
  uint32_t tile_region;     // bytes of the tile region (the forest chunk follows it)
  float* preds;             // partial sums between chunk passes (local scratch; may alias `out`)
  float* out;               // final scores: written once, by the last chunk pass -- may be a PEER-mapped buffer
                            // (rank g's slice of rank 0's gathered output): never read back
  int32_t* leaf_ids;
  const int32_t* leaf_origin;  // [T][2^d]
  Transform tf;
  // k_smem_forest_deep (sparse forests, QT layout): the trees in qwords are the dense tops; a bottom slot that
  // is a continuation marker parks the walk, which is finished in L2 on the rank-form breadth-first nodes
  const uint4* deep_blocks;      // [blocks + 1][2] two-level blocks {root, left, right, -} x {rank word | leaf value, leaf<<31 | block delta}
  const int64_t* block_offsets;  // [T+1] first block of every tree
  const int32_t* block_orig;     // [blocks + 1][3] caller's node index of root / left / right
};

// Levels 0 and 1 of every tree of a launch -- {root, left child, right child}, 12 bytes per tree -- travel as a
// KERNEL PARAMETER (constant bank 0; parameters may be up to 32,764 bytes since CUDA 12.1).  The root is
// the same word for all lanes and the level-1 node is a select between two warp-uniform words, so both
// come through the uniform datapath (LDCU / ULDC) instead of two shared-memory loads: two of a tree's
// ~28-38 LSU wavefronts and two of its DEPTH + 1 dependent shared-memory round trips disappear.  A forest
// with more trees than the table holds is run as several launches over consecutive chunk ranges
// (SmemParams::tree_base / first_pass / last_pass), each with the table of its own trees.
// (kTopTrees, TopWords: forest.h)

constexpr uint32_t kContMask = 0xFFC00000u;   // packer.cu: marker = kContMask | node index relative to the chunk's first node

constexpr int kSmemMaxThreads = kSmemMaxWarps * 32;

// Shared-memory map of k_smem_forest (byte offsets from smem[]):
//   [0, tile_region)               tile region: word (col, slot, lane) at col*col_stride + slot*slot_stride
//                                  + lane*4 -- the region base is one CTA-uniform constant and a
//                                  lane's address is a pure bit-field OR
//   [.., + trees_per_chunk*W*4)    the resident chunk of rank-quantised trees
//   [.., + (1 + slots)*8)          mbarriers: chunk, then one per tile buffer
// slots = warps * 2 (tiles double-buffered by TMA bulk copies) or warps * 1 (STAGED: the next tile
// waits in registers while the current one is traversed -- half the tile memory, so deep trees get
// more of them per chunk).

// One group of G trees for this lane's sample, on absolute shared-memory addresses.
//   tree_abs   shared address of word 0 of the group's first tree (warp-uniform)
//   s_base     shared address of smem[0] = base of the tile region (CTA-uniform)
//   slot_lane  slot*128 + lane*4 : this lane's bank inside this warp's current tile buffer
// a[g] is the shared ADDRESS of the current node, so one visit is six instructions:
//   LDS  node  [a]
//   LOP3 (node & col_mask) | slot_lane
//   LDS  rank  [that + s_base]                   (uniform-register base, folded into the LDS)
//   ISETP rank > node ; SEL k = right ? 4-B : -B ; IADD3 a = a + a + k      (B = tree address)
// i.e. a' - B = 2(a - B) + 4*right, the heap step on byte offsets, with no separate address add.
// Loads are asm so the address arithmetic stays exactly this; ptxas still interleaves the G chains.
template <int DEPTH, int G, bool MISSING, int TOPL>
__device__ __forceinline__ void traverse_group(uint32_t tree_abs, uint32_t s_base, uint32_t slot_lane,
                                               uint32_t col_mask, uint32_t miss_code, uint32_t (&a)[G],
                                               const TopWords& tw, int top_index) {
  constexpr uint32_t W4 = 4u << (DEPTH + 1);  // bytes per tree
  uint32_t k0[G];
#pragma unroll
  for (int g = 0; g < G; ++g) k0[g] = 0u - (tree_abs + g * W4);
  auto right = [&](uint32_t F, uint32_t N) {
    bool c = F > N;
    if (MISSING) {
      if (F == miss_code) c = !(N & 1u);      // sentinel: !def_left (BaseTahoeTest.h:452)
    }
    return c;
  };
  // TOPL = levels decided on parameter words (constant bank): 0 = none, 1 = the root, 2 = root + level 1
  constexpr int kTop = TOPL >= 2 && DEPTH >= 2 ? 2 : (TOPL >= 1 ? 1 : 0);
  if (kTop > 0) {
    uint32_t R0[G], F0[G];
    bool c0[G];
#pragma unroll
    for (int g = 0; g < G; ++g) R0[g] = tw.w[3 * (top_index + g)];
#pragma unroll
    for (int g = 0; g < G; ++g)   // root: warp-uniform word, the column offset is uniform arithmetic
      F0[g] = lds_u32(s_base + ((R0[g] & col_mask) | slot_lane));
#pragma unroll
    for (int g = 0; g < G; ++g) c0[g] = right(F0[g], R0[g]);
    if (kTop == 2) {
      // level 1: a select between two warp-uniform words (ptxas makes it a pair of predicated constant
      // loads behind the compare: a win with 12+ warps to cover them -- config 2 8.28 -> 8.10 ms, config 4
      // 9.00 -> 8.84 -- a loss with 8: config 3 on 2M rows 11.57 -> 12.00 ms; the launcher decides)
      uint32_t N1[G], F1[G];
#pragma unroll
      for (int g = 0; g < G; ++g)
        N1[g] = c0[g] ? tw.w[3 * (top_index + g) + 2] : tw.w[3 * (top_index + g) + 1];
#pragma unroll
      for (int g = 0; g < G; ++g) F1[g] = lds_u32(s_base + ((N1[g] & col_mask) | slot_lane));
#pragma unroll
      for (int g = 0; g < G; ++g)   // heap index at level 2 = 4 + 2 c0 + c1 (root = 1)
        a[g] = tree_abs + g * W4 + 16u + (c0[g] ? 8u : 0u) + (right(F1[g], N1[g]) ? 4u : 0u);
    } else {
#pragma unroll
      for (int g = 0; g < G; ++g) a[g] = tree_abs + g * W4 + 8u + (c0[g] ? 4u : 0u);
    }
  } else {
#pragma unroll
    for (int g = 0; g < G; ++g) a[g] = tree_abs + g * W4 + 4;
  }
#pragma unroll
  for (int l = kTop; l < DEPTH; ++l) {
    uint32_t N[G], F[G];
#pragma unroll
    for (int g = 0; g < G; ++g) N[g] = lds_u32(a[g]);
#pragma unroll
    for (int g = 0; g < G; ++g) {
      F[g] = lds_u32(s_base + ((N[g] & col_mask) | slot_lane));
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const bool c = right(F[g], N[g]);
      a[g] = a[g] + a[g] + (c ? k0[g] + 4u : k0[g]);
    }
  }
}

__device__ __forceinline__ void sts_u128(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// SINGLE: one tile buffer per warp filled by TMA (no register staging, no second buffer): the warp asks for its
// next tile into L2 (cp.async.bulk.prefetch.L2) when it starts to walk and issues the bulk copy after the walk, so
// the copy reads L2 and the SM's other warps cover that short wait.  The staged plans pay 2 x cols / 4 128-bit LSU
// instructions per lane and tile for the same saving of shared memory.
template <int DEPTH, int G, bool LEAF_IDS, bool STAGED, int TOPL, bool SINGLE = false>
__global__ void __launch_bounds__(kSmemMaxThreads, 1)
k_smem_forest(const __grid_constant__ SmemParams p, const __grid_constant__ TopWords tw) {
  static_assert(!LEAF_IDS || TOPL == 0, "the leaf-id instantiation keeps every level in shared memory");
  static_assert(!(STAGED && SINGLE), "one way of filling the only buffer");
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr uint32_t W = 1u << (DEPTH + 1);      // words per tree
  constexpr uint32_t L = 1u << DEPTH;            // bottom slots per tree
  constexpr int kBufs = (STAGED || SINGLE) ? 1 : 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  const int slots = nwarps * kBufs;
  const uint32_t tile_bytes = (uint32_t)p.cols * 128u;
  const uint32_t forest_off = p.tile_region;
  const uint32_t bars_off = forest_off + (uint32_t)p.trees_per_chunk * W * 4u;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t bar_chunk = s_base + bars_off;
  const uint32_t bar_tile0 = bar_chunk + 8 + warp * (8 * kBufs);

  if (threadIdx.x == 0) {
    mbar_init(bar_chunk, 1);
    for (int i = 0; i < slots; ++i) mbar_init(bar_chunk + 8 + i * 8, 1);
    fence_mbar_init();
  }
  __syncthreads();

  const size_t warp_global = (size_t)blockIdx.x * nwarps + warp;
  const size_t warp_stride = (size_t)gridDim.x * nwarps;
  uint32_t tile_phase = 0;   // bit b = parity to wait for on this warp's buffer b
  uint32_t chunk_phase = 0;

  // TMA path: bring tile `tile` into this warp's buffer `b`.  Slot-major region (col_stride == 128):
  // the buffer is the tile's own [col][32 lanes] image, so ONE bulk copy of cols*128 bytes brings it in
  // (one UBLKCP and one proxy fence per tile instead of one per column; the per-column form made the
  // TMA path lose to register staging on 28-column tiles, profiles/r01_smem_plans.log).  Column-major
  // region: the slots of a column are interleaved, one 128-byte copy per column row.
  const bool tile_contiguous = p.col_stride == 128u;
  auto load_tile = [&](size_t tile, int b) {
    const uint32_t bar = bar_tile0 + b * 8;
    const uint32_t slot_off = (uint32_t)(warp * kBufs + b) * p.slot_stride;
    const uint32_t* src = p.qtiles + tile * (size_t)p.cols * 32;
    if (tile_contiguous) {
      if (lane == 0) {
        fence_proxy_async();
        mbar_expect_tx(bar, tile_bytes);
        bulk_g2s(s_base + slot_off, src, tile_bytes, bar);
      }
      return;
    }
    if (lane == 0) mbar_expect_tx(bar, tile_bytes);
    __syncwarp();
    for (int c = lane; c < p.cols; c += 32) {
      fence_proxy_async();
      bulk_g2s(s_base + c * p.col_stride + slot_off, src + c * 32, 128u, bar);
    }
  };
  // STAGED path: the tile is cols*8 consecutive 16-byte pieces in global memory ([col][32 lanes]);
  // lane l keeps pieces l, l+32, ... in registers and stores each to its column row
  constexpr int kPieces = kStagedMaxCols / 4;            // per lane
  uint4 st[STAGED ? kPieces : 1];
  const int npieces = p.cols * 8;
  auto fetch_tile = [&](size_t tile) {
    const uint4* src = reinterpret_cast<const uint4*>(p.qtiles + tile * (size_t)p.cols * 32) + lane;
#pragma unroll
    for (int j = 0; j < (STAGED ? kPieces : 0); ++j)
      if (lane + 32 * j < npieces) st[j] = __ldg(src + 32 * j);
  };
  auto commit_tile = [&]() {
    // piece i = lane + 32 j: column i / 8, bytes (i % 8) * 16 of that column's 128-byte row
    const uint32_t dst = s_base + (uint32_t)warp * p.slot_stride + (uint32_t)(lane & 7) * 16u +
                         (uint32_t)(lane >> 3) * p.col_stride;
#pragma unroll
    for (int j = 0; j < (STAGED ? kPieces : 0); ++j)
      if (lane + 32 * j < npieces) sts_u128(dst + (uint32_t)(4 * j) * p.col_stride, st[j]);
  };

  for (int chunk = 0; chunk < p.num_chunks; ++chunk) {
    const int tree0 = chunk * p.trees_per_chunk;
    const int ntrees = min(p.trees_per_chunk, p.padded_trees - tree0);   // a multiple of G
    // ---- stage this chunk of trees (everyone is done reading the previous one)
    __syncthreads();
    if (threadIdx.x == 0) {
      fence_proxy_async();
      const uint32_t bytes = (uint32_t)ntrees * W * 4u;
      mbar_expect_tx(bar_chunk, bytes);
      const uint8_t* src = (const uint8_t*)(p.qwords + (size_t)tree0 * W);
      uint32_t done = 0;
      while (done < bytes) {               // pieces of <= 64 KB
        const uint32_t piece = min(bytes - done, 65536u);
        bulk_g2s(s_base + forest_off + done, src + done, piece, bar_chunk);
        done += piece;
      }
    }
    // the first tile of this pass flies while the chunk lands
    size_t tile = warp_global;
    if (tile < p.num_tiles) {
      if (STAGED) fetch_tile(tile); else load_tile(tile, 0);
    }
    mbar_wait(bar_chunk, chunk_phase);
    chunk_phase ^= 1;

    const bool first = p.first_pass && chunk == 0, last = p.last_pass && chunk == p.num_chunks - 1;
    int buf = 0;
    // the tile's missing flag and this lane's partial sum are fetched one tile ahead, so that their
    // global-memory latency hides behind the traversal of the tile before
    uint32_t flag_next = 0;
    float acc_next = 0.0f;
    if (tile < p.num_tiles) {
      flag_next = __ldg(p.tile_flags + tile);
      const size_t r0 = tile * 32 + lane;
      if (!first && r0 < p.rows) acc_next = p.preds[r0];
    }
    for (; tile < p.num_tiles; tile += warp_stride) {
      const size_t next = tile + warp_stride;
      __syncwarp();                          // all lanes finished reading the buffer about to be refilled
      const bool has_missing = flag_next != 0;
      float acc = acc_next;
      if (STAGED) {
        commit_tile();                       // registers -> this warp's only buffer
        __syncwarp();
        if (next < p.num_tiles) fetch_tile(next);   // lands while this tile is traversed
      } else if (SINGLE) {
        if (next < p.num_tiles && lane == 0)          // HBM -> L2 under this tile's walk
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.qtiles + next * (size_t)p.cols * 32),
                       "r"(tile_bytes) : "memory");
      } else if (next < p.num_tiles) {
        load_tile(next, buf ^ 1);
      }
      if (next < p.num_tiles) {
        flag_next = __ldg(p.tile_flags + next);
        const size_t rn = next * 32 + lane;
        acc_next = (!first && rn < p.rows) ? p.preds[rn] : 0.0f;
      }
      if (!STAGED) {
        mbar_wait(bar_tile0 + buf * 8, (tile_phase >> buf) & 1u);
        tile_phase ^= 1u << buf;
      }

      const size_t row = tile * 32 + lane;
      const bool live = row < p.rows;
      const uint32_t slot_lane = (uint32_t)(warp * kBufs + buf) * p.slot_stride + lane * 4u;

      for (int t = 0; t < ntrees; t += G) {
        uint32_t a[G];
        const uint32_t tree_abs = s_base + forest_off + (uint32_t)t * W * 4u;
        // TOPL is a template parameter of the KERNEL: a run-time choice between the forms puts every form's
        // unrolled walk into one kernel, and the walk then no longer fits the instruction cache (measured:
        // config 3, 8 warps, 11.57 -> 12.1 ms with three forms inlined even when the old one is the one run)
        if (has_missing) traverse_group<DEPTH, G, true, TOPL>(tree_abs, s_base, slot_lane, p.col_mask, p.miss_code, a, tw, tree0 + t);
        else traverse_group<DEPTH, G, false, TOPL>(tree_abs, s_base, slot_lane, p.col_mask, p.miss_code, a, tw, tree0 + t);
#pragma unroll
        for (int g = 0; g < G; ++g) {
          acc += __uint_as_float(lds_u32(a[g]));   // a padding tree adds +0.0f: the sum keeps its bits
          // (the leaf-id instantiation is never split over launches: tree_base == 0)
          if (LEAF_IDS && live && tree0 + t + g < p.num_trees)
            p.leaf_ids[row * (size_t)p.num_trees + tree0 + t + g] =
                __ldg(p.leaf_origin + (size_t)(tree0 + t + g) * L + ((a[g] - tree_abs - g * W * 4u) / 4 - L));
        }
      }
      if (live) {
        if (last) p.out[row] = apply_transform(acc, p.tf);
        else p.preds[row] = acc;
      }
      if (SINGLE) {
        __syncwarp();                        // every lane is done with the buffer
        if (next < p.num_tiles) load_tile(next, 0);
      } else if (!STAGED) {
        buf ^= 1;
      }
    }
  }
}

}  // namespace tahoe
