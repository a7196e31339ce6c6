// kernels_smem_deep.cuh -- k_smem_forest_deep: SHARED_FOREST / SPLIT_SHARED_FOREST on SPARSE forests
// (QT layout, packer.cu plan_sparse_qtop).  The reference's sparse forest is disabled
// (BaseTahoeTest.h:727-846); its walk is Struct.h:2231-2244, its shared-forest strategies are
// Struct.h:1245-1606.  Included by kernels_smem_depth.cu (one translation unit per top depth).
//
// The top DEPTH levels of every tree are a dense rank-quantised heap in shared memory, walked by
// the six-instruction visit of k_smem_forest (traverse_group).  A walk that is still on an internal
// node at the bottom of the top reads a continuation marker there and is PARKED in a lane-private
// ring; it is finished in L2 on two-level blocks of 32 bytes {root, left, right, -} in rank form
// (`deep_blocks`): one sector decides two levels.
//
// The L2 round trips of those deep walks never stall the warp: every lane keeps K deep walks in
// flight and advances each by ONE block per traversal group -- the block loads (LDG.256) are issued
// before the group's shared-memory walk (~1.5 k cycles) and consumed after it.  So that a deep walk may outlive
// the tile that parked it, a warp owns TWO tile buffers (current and previous tile; the next tile
// waits in registers): a walk reads its features from the buffer of its own tile, the score of tile
// n is written at the end of tile n + 1, and only the walks of tile n - 1 that are STILL unfinished
// then (few: they had a whole tile of groups to run) are finished with the latency exposed.
#pragma once
#include "kernels_smem_traverse.cuh"

namespace tahoe {

// one 32-byte block, read-only (LDG.E.256); volatile so that it is issued where it is written (before
// the shared-memory walk that hides its latency), not sunk to its first use
__device__ __forceinline__ void ldg_nc_block(uint64_t addr, uint4& a, uint4& b) {
  asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
               : "l"(addr));
}

// PAIR: rows wider than 32 columns need 64 staging registers per lane, which caps a CTA at 8 warps --
// two per scheduler, too few to cover the fixed-latency dependencies of the walk.  PAIR runs 16 warps
// with the shared memory of 8: the two warps of a pair work on the SAME tile (each stages half of its
// pieces and walks every other traversal group of the chunk), meet at a named barrier before and after
// the tile is committed, and warp 1 hands its part of a row's sum to warp 0 through shared memory
// (fixed order: the score stays reproducible).
//
// Shared-memory map (byte offsets from smem[]):
//   [0, tile_region)               tile region, TWO buffers per unit (unit = warp, or pair of warps;
//                                  slot = unit*2 + parity of the tile)
//   [.., + trees_per_chunk*W*4)    the resident chunk of dense tops
//   [.., + 512)                    the chunk mbarrier
//   [.., + warps*kDeepPend*128)    parked walks: lane-private rings, entry i of lane l at (i*32 + l)*4
//   [.., + 1024)                   PAIR: one float per lane per pair, warp 1 -> warp 0
// PIECES: 16-byte pieces of the staged (next) tile one lane keeps in registers (8: <= 32 columns, or
// <= 64 as a pair; 16: <= 64 columns, plans of <= 8 warps).  K: deep walks in flight per lane.
template <int DEPTH, int G, bool LEAF_IDS, int PIECES, int K, bool PAIR>
__global__ void __launch_bounds__(PIECES > kStagedMaxCols / 4 ? 256 : kSmemMaxThreads, 1)
k_smem_forest_deep(const __grid_constant__ SmemParams p, const __grid_constant__ TopWords tw) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr uint32_t W = 1u << (DEPTH + 1);      // words per tree
  constexpr uint32_t L = 1u << DEPTH;            // bottom slots per tree
  constexpr uint32_t kOffMask = kDeepPend * 128 - 1;   // ring positions are byte offsets, 128 apart
  static_assert((kDeepPend & (kDeepPend - 1)) == 0, "ring size must be a power of two");
  static_assert(kDeepPend > kMaxGroup + 1, "a traversal group must fit the ring");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  const int unit = PAIR ? warp >> 1 : warp, half = PAIR ? warp & 1 : 0;
  const int nunits = PAIR ? nwarps >> 1 : nwarps;
  constexpr int kSplit = PAIR ? 2 : 1;           // warps that share a tile
  const uint32_t forest_off = p.tile_region;
  const uint32_t bars_off = forest_off + (uint32_t)p.trees_per_chunk * W * 4u;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t bar_chunk = s_base + bars_off;
  uint8_t* ring = smem + bars_off + 512 + warp * (kDeepPend * 128) + lane * 4;   // entry at ring + offset
  float* xch = reinterpret_cast<float*>(smem + bars_off + 512 + nwarps * (kDeepPend * 128)) + unit * 32 + lane;
  // the two warps of a pair meet here (named barrier 1 + pair); one warp alone only re-converges
  auto unit_sync = [&]() {
    if (PAIR) asm volatile("bar.sync %0, 64;" ::"r"(1 + unit) : "memory");
    else __syncwarp();
  };

  if (threadIdx.x == 0) {
    mbar_init(bar_chunk, 1);
    fence_mbar_init();
  }
  __syncthreads();

  const size_t warp_global = (size_t)blockIdx.x * nunits + unit;   // tiles are dealt to units
  const size_t warp_stride = (size_t)gridDim.x * nunits;
  uint32_t chunk_phase = 0;

  // the next tile waits in registers: cols*8 consecutive 16-byte pieces of [col][32 lanes]; this warp
  // keeps pieces lane + 32 * (kSplit * j + half)
  uint4 st[PIECES];
  const int npieces = p.cols * 8;
  auto fetch_tile = [&](size_t tile) {
    const uint4* src = reinterpret_cast<const uint4*>(p.qtiles + tile * (size_t)p.cols * 32) + lane;
#pragma unroll
    for (int j = 0; j < PIECES; ++j)
      if (lane + 32 * (kSplit * j + half) < npieces) st[j] = __ldg(src + 32 * (kSplit * j + half));
  };
  auto commit_tile = [&](int b) {
    // piece i: column i / 8, bytes (i % 8) * 16 of that column's 128-byte row
    const uint32_t dst = s_base + (uint32_t)(unit * 2 + b) * p.slot_stride + (uint32_t)(lane & 7) * 16u +
                         (uint32_t)(lane >> 3) * p.col_stride;
#pragma unroll
    for (int j = 0; j < PIECES; ++j)
      if (lane + 32 * (kSplit * j + half) < npieces)
        sts_u128(dst + (uint32_t)(4 * (kSplit * j + half)) * p.col_stride, st[j]);
  };

  // ---- deep walks of this lane -----------------------------------------------------------
  // A slot always walks something: an idle slot sits on the DUMMY leaf the packer appends after the
  // last node (value +0.0f), so a step needs no "slot in use" state -- reaching a leaf, real or
  // dummy, adds its value and frees the slot for the next parked walk.
  const uint32_t sl0 = (uint32_t)(unit * 2) * p.slot_stride + lane * 4u;   // this lane in the unit's buffer 0
  uint32_t node[K];            // block index, relative to the chunk's first block
  uint32_t fb[K];              // sl0 or sl0 + slot_stride: the tile buffer the walk's row lives in
  uint4 ea[K], eb[K];          // block `node[w]`, loaded by the previous step: {root.x, root.y, left.x, left.y}, {right.x, right.y, -, -}
#pragma unroll
  for (int w = 0; w < K; ++w) {
    node[w] = 0; fb[w] = sl0;
    ea[w] = make_uint4(0u, kMetaLeaf, 0u, kMetaLeaf); eb[w] = make_uint4(0u, kMetaLeaf, 0u, 0u);
  }
  // ring of parked walks {node index | tile parity << 31}: pop at head, push at tail (byte offsets;
  // at most kDeepPend - 1 entries, so head == tail means empty)
  uint32_t head = 0, tail = 0;

  for (int chunk = 0; chunk < p.num_chunks; ++chunk) {
    const int tree0 = chunk * p.trees_per_chunk;
    const int ntrees = min(p.trees_per_chunk, p.padded_trees - tree0);   // a multiple of G
    __syncthreads();                         // everyone is done reading the previous chunk
    if (threadIdx.x == 0) {
      fence_proxy_async();
      const uint32_t bytes = (uint32_t)ntrees * W * 4u;
      mbar_expect_tx(bar_chunk, bytes);
      const uint8_t* src = (const uint8_t*)(p.qwords + (size_t)tree0 * W);
      uint32_t done = 0;
      while (done < bytes) {
        const uint32_t piece = min(bytes - done, 65536u);
        bulk_g2s(s_base + forest_off + done, src + done, piece, bar_chunk);
        done += piece;
      }
    }
    size_t tile = warp_global;
    if (tile < p.num_tiles) fetch_tile(tile);
    mbar_wait(bar_chunk, chunk_phase);
    chunk_phase ^= 1;

    const bool first = p.first_pass && chunk == 0, last = p.last_pass && chunk == p.num_chunks - 1;
    const int64_t chunk_first = __ldg(p.block_offsets + min(p.tree_base + tree0, p.num_trees));
    const uint32_t dummy = (uint32_t)(__ldg(p.block_offsets + p.num_trees) - chunk_first);
    uint64_t cdv = (uint64_t)(p.deep_blocks + 2 * chunk_first);   // kept as one 64-bit register: block address = one IMAD.WIDE
    asm volatile("" : "+l"(cdv));
#pragma unroll
    for (int w = 0; w < K; ++w) node[w] = dummy;       // nothing is in flight at a chunk boundary
    int buf = 0;
    // the tile's missing flag and this lane's partial sum are fetched one tile ahead
    uint32_t flag_next = 0;
    float acc_next = 0.0f;
    if (tile < p.num_tiles) {
      flag_next = __ldg(p.tile_flags + tile);
      const size_t r0 = tile * 32 + lane;
      if (!first && half == 0 && r0 < p.rows) acc_next = p.preds[r0];
    }

    // current / previous tile of this warp
    uint32_t cur_sl = sl0, prev_sl = 0xFFFFFFFFu, prev_par = 0;
    size_t row = 0, row_prev = 0;
    bool have_prev = false, live_prev = false;
    float acc_prev = 0.0f;       // top part of the previous tile's sum
    float dacc_prev = 0.0f, dacc_cur = 0.0f;   // leaf values the deep walks of each tile have reached
    // PAIR, warp 0: its part of a retired tile's sum, written once warp 1's part has crossed the barrier
    float out_sum = 0.0f;
    size_t out_row = 0;
    bool out_have = false;
    auto flush_out = [&]() {
      if (PAIR && half == 0 && out_have) {
        const float s = out_sum + *xch;
        if (last) p.out[out_row] = apply_transform(s, p.tf);
        else p.preds[out_row] = s;
        out_have = false;
      }
    };

    // two levels of every walk in flight: consume the block loaded by the previous step (root, then the
    // child it picks), start a parked walk in a slot that reached its leaf, issue the next block load
    auto deep_step = [&]() {
#pragma unroll
      for (int w = 0; w < K; ++w) {
        const uint4 a = ea[w], b = eb[w];
        const bool root_leaf = (int32_t)a.y < 0;
        // (a leaf's x is its float value: keep its bits out of the address)
        const uint32_t F0 = lds_u32(s_base + (((root_leaf ? 0u : a.x) & p.col_mask) | fb[w]));
        bool c0 = F0 > a.x;
        if (F0 == p.miss_code) c0 = !(a.x & 1u);                 // sentinel: !def_left (BaseTahoeTest.h:452)
        const uint32_t cx = c0 ? b.x : a.z, cy = c0 ? b.y : a.w; // the child the root picks
        const bool done = root_leaf || (int32_t)cy < 0;
        const uint32_t F1 = lds_u32(s_base + (((done ? 0u : cx) & p.col_mask) | fb[w]));
        bool c1 = F1 > cx;
        if (F1 == p.miss_code) c1 = !(cx & 1u);
        const float v = done ? __uint_as_float(root_leaf ? a.x : cx) : 0.0f;
        const bool pv = fb[w] == prev_sl;
        if (pv) dacc_prev += v; else dacc_cur += v;
        if (LEAF_IDS && done && node[w] != dummy) {
          const int64_t blk = chunk_first + (int64_t)node[w];
          int lo = tree0, hi = min(tree0 + ntrees, p.num_trees);   // block_offsets[lo] <= blk < block_offsets[hi]
          while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(p.block_offsets + mid) <= blk) lo = mid; else hi = mid;
          }
          p.leaf_ids[(pv ? row_prev : row) * (size_t)p.num_trees + lo] =
              __ldg(p.block_orig + 3 * blk + (root_leaf ? 0 : (c0 ? 2 : 1)));
        }
        const bool can = done && head != tail;
        const uint32_t ent = *reinterpret_cast<const uint32_t*>(ring + head);   // (stale when !can: unused)
        head = can ? (head + 128u) & kOffMask : head;
        const uint32_t stepped = node[w] + cy + (c1 ? 1u : 0u);   // (!done: cy is the block delta)
        node[w] = can ? (ent & 0x7FFFFFFFu) : (done ? dummy : stepped);
        fb[w] = can ? sl0 + (ent >> 31) * p.slot_stride : fb[w];
        ldg_nc_block(cdv + (uint64_t)node[w] * 32u, ea[w], eb[w]);
      }
    };
    // finish what is left of the previous tile's walks (latency exposed), then write its score
    auto retire_prev = [&]() {
      while (true) {
        // the ring is first-in first-out: entries of the previous tile, if any, are at its head
        bool mine = head != tail && (*reinterpret_cast<const uint32_t*>(ring + head) >> 31) == prev_par;
#pragma unroll
        for (int w = 0; w < K; ++w) mine |= node[w] != dummy && fb[w] == prev_sl;
        if (!__any_sync(0xffffffffu, mine)) break;
        deep_step();
      }
      const float s = acc_prev + dacc_prev;
      if (!PAIR) {
        if (live_prev) {
          if (last) p.out[row_prev] = apply_transform(s, p.tf);
          else p.preds[row_prev] = s;
        }
      } else if (half == 1) {
        *xch = s;                                // read by warp 0 after the pair's next barrier
      } else {
        out_sum = s; out_row = row_prev; out_have = live_prev;
      }
    };

    for (; tile < p.num_tiles; tile += warp_stride) {
      const size_t next = tile + warp_stride;
      unit_sync();                           // buffer `buf` held tile n-2: all of its walks are retired
      flush_out();
      float acc = acc_next;
      const bool has_missing = flag_next != 0;
      commit_tile(buf);
      unit_sync();
      if (next < p.num_tiles) {
        fetch_tile(next);                    // lands while this tile is traversed
        flag_next = __ldg(p.tile_flags + next);
        const size_t rn = next * 32 + lane;
        acc_next = (!first && half == 0 && rn < p.rows) ? p.preds[rn] : 0.0f;
      }
      row = tile * 32 + lane;
      const bool live = row < p.rows;
      cur_sl = sl0 + (uint32_t)buf * p.slot_stride;
      const uint32_t park_tag = (uint32_t)buf << 31;
      const bool park = live;

      for (int t = half * G; t < ntrees; t += kSplit * G) {
        // room for G more parked walks in every lane's ring (rarely taken)
        while (__any_sync(0xffffffffu, ((tail - head) & kOffMask) > (uint32_t)(kDeepPend - 1 - G) * 128u)) deep_step();
        deep_step();
        uint32_t a[G];
        const uint32_t tree_abs = s_base + forest_off + (uint32_t)t * W * 4u;
        constexpr int TOPL = LEAF_IDS ? 0 : 2;       // 16 warps (or pairs): both parameter levels pay (config 4: 9.00 -> 8.84 ms)
        if (has_missing) traverse_group<DEPTH, G, true, TOPL>(tree_abs, s_base, cur_sl, p.col_mask, p.miss_code, a, tw, tree0 + t);
        else traverse_group<DEPTH, G, false, TOPL>(tree_abs, s_base, cur_sl, p.col_mask, p.miss_code, a, tw, tree0 + t);
        uint32_t v[G];
#pragma unroll
        for (int g = 0; g < G; ++g) v[g] = lds_u32(a[g]);
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const bool cont = park && v[g] >= kContMask;           // marker: all of the top ten bits set
          acc += cont ? 0.0f : __uint_as_float(v[g]);
          // the slot at tail is free: write it anyway, keep it only for a parked walk
          *reinterpret_cast<uint32_t*>(ring + tail) = (v[g] & ~kContMask) | park_tag;
          tail = cont ? (tail + 128u) & kOffMask : tail;
          if (LEAF_IDS && live && !cont && tree0 + t + g < p.num_trees)
            p.leaf_ids[row * (size_t)p.num_trees + tree0 + t + g] =
                __ldg(p.leaf_origin + (size_t)(tree0 + t + g) * L + ((a[g] - tree_abs - g * W * 4u) / 4 - L));
        }
      }
      if (have_prev) retire_prev();
      // this tile becomes the previous one
      have_prev = true; live_prev = live; row_prev = row; prev_sl = cur_sl;
      acc_prev = acc; dacc_prev = dacc_cur; dacc_cur = 0.0f;
      prev_par = (uint32_t)buf;
      buf ^= 1;
    }
    // end of the chunk: nothing stays in flight
    if (PAIR) {
      unit_sync();
      flush_out();                           // the tile before the last one
      unit_sync();
    }
    if (have_prev) retire_prev();
    if (PAIR) {
      unit_sync();
      flush_out();
    }
  }
}

}  // namespace tahoe
