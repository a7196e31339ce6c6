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
//
// Round 2 (profiles/r02_c1_walk_cycles.md): wide rows run as two-warp teams with ONE row buffer each; the team's
// next row is prefetched into L2 while the current one is walked (cp.async.bulk.prefetch.L2), the forest's top
// levels are one bulk-copied image without padding slots, and the blocked walk reads nodes stored the way a
// decision uses them (kSd* in forest.h; sd_child).
#include <cstdlib>
#include <map>
#include <mutex>
#include <utility>

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
  int teams_per_block;       // a team = team_warps warps sharing one row (row_bufs row buffers)
  int team_warps;            // 1, 2 or 4
  int top_levels;            // levels 0..top_levels-1 of the interleaved forest are copied to shared memory
  uint32_t top_off;          // byte offset of that copy ([2^top_levels - 1][top_stride] uint2)
  int top_stride;            //   slots per node in it: tpad, or num_trees (blocked walk: padding slots left out)
  uint32_t top_bytes;        //   its size (blocked walk: the packer's image, a multiple of 16 bytes, copied as it is)
  // B32 (BLOCKED kernels): the copy comes from top_nodes (leaves pushed down); levels top_levels..depth
  // are walked two at a time through 32-byte blocks {root, left, right, -}
  const uint2* top_nodes;    // [2^top_levels - 1][top_stride]: the image of the shared-memory copy
  const uint4* blocks;       // [block][tpad][2]
  const int32_t* orig_bottom;  // [2^depth][tpad]
  int block_levels;          // L2 trips below the top: (depth + 1 - top_levels) / 2 >= 1
  int l2_ahead;              // single-buffer teams: prefetch the next row into L2 during the walk
  int row_bufs;              // row buffers per team: 2 = the next row streams in under the walk; 1 = twice the teams,
                             // a team's own load latency exposed, hidden by the other teams of the SM
};

__device__ __forceinline__ void ldg_block(const uint4* p, uint32_t (&w)[8]) {   // one 32-byte sector, one request
  asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
               : "l"(p));
}


__device__ __forceinline__ void team_barrier(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// One row, this warp's share of the tree slots (slot groups wt, wt + team_warps, ...), G at a time
// (G = tree slots walked together by one lane: ILP over the L2 latency).  Returns the warp's partial sum (double: lane = tree changes the summation order, so the
// sum is kept within half an ulp of exact).  The level step is branch-free -- a walk that has
// reached its leaf re-reads that leaf -- so the kSdGroup dependent chains (L2 gather -> feature ->
// compare -> next index) interleave in one warp.
template <int kSdGroup>
__device__ __forceinline__ double sd_walk_row(const SharedDataParams& p, const uint2* s_top, const float* my,
                                              size_t row, int lane, int wt) {
  double acc = 0.0;
  const int gstride = 32 * p.team_warps;
  for (int s0 = lane + 32 * wt; s0 < p.tpad; s0 += gstride * kSdGroup) {
    uint32_t n[kSdGroup];
    uint2 e[kSdGroup];
    const uint2* col[kSdGroup];
    const uint2* top[kSdGroup];
#pragma unroll
    for (int g = 0; g < kSdGroup; ++g) {
      const int slot = min(s0 + gstride * g, p.tpad - 1);   // past the end: repeat the last slot, drop its sum
      col[g] = p.inter + slot;
      top[g] = s_top + slot;
      n[g] = 0;
      e[g] = p.top_levels > 0 ? top[g][0] : __ldg(col[g]);
    }
    // one level: feature -> compare -> child index (a walk on its leaf stays there)
    auto step = [&](int g) -> bool {
      const bool leaf = (e[g].y & kMetaLeaf) != 0;
      const float v = my[leaf ? 0u : (e[g].y & kMetaFidMask)];
      bool c = branch_right(v, __uint_as_float(e[g].x), e[g].y & kMetaDefLeft, p.missing);
      c = c != ((e[g].y & kMetaExchange) != 0);       // Struct.h:980
      n[g] = leaf ? n[g] : 2 * n[g] + 1 + (c ? 1u : 0u);
      return !leaf;
    };
    // The top levels come from shared memory (bank = slot: conflict-free); once the lanes' paths
    // have parted, every lane of a global load sits in another sector, and each of those L2 round
    // trips costs ~800 cycles of a dependent chain -- so only the last levels pay it.
    int l = 0;
    for (; l < min(p.depth, p.top_levels - 1); ++l) {
#pragma unroll
      for (int g = 0; g < kSdGroup; ++g) {
        step(g);
        e[g] = top[g][(size_t)n[g] * p.tpad];
      }
    }
    for (; l < p.depth; ++l) {
      bool live = false;
#pragma unroll
      for (int g = 0; g < kSdGroup; ++g) {
        live |= step(g);
        e[g] = __ldg(col[g] + (size_t)n[g] * p.tpad);
      }
      if (!__any_sync(0xffffffffu, live)) break;   // every walk of the warp already sat on a leaf
    }
#pragma unroll
    for (int g = 0; g < kSdGroup; ++g) {
      const int slot = s0 + gstride * g;
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

// B32 walk of one row: top_levels (>= 1) levels from shared memory, then block_levels L2 trips of
// two levels each.  Leaves are pushed down in this layout, so every walk makes exactly `depth`
// decisions and ends on the bottom level -- no leaf test in the loop, no early exit.
// The kernel is bound by the INSTRUCTIONS of these walks (clock64 around the walk: its time grows in
// proportion to the teams per SM whatever the split between shared-memory levels and L2 trips;
// profiles/r02_c1_walk_cycles.md), so the node's second word is stored the way a decision uses it
// (kSd* in forest.h) and all addresses are 32-bit byte offsets.
//   top / my: the top levels [node][tpad] uint2 and the team's row, both in shared memory.
// The child SLOT a walk takes at one node, as 1 / 2 (the heap step is n = 2n + that): the reference's test (BaseTahoeTest.h:450-452: a missing value
// follows the default direction, else v >= thr goes right) with the exchange of Struct.h:980 folded into the two
// flag bits of the node word.  Written as PTX so that each flag costs one predicate-setting instruction.
// NANM: the sentinel is NaN, so "missing" means "is NaN" (device_utils.cuh, is_missing).
template <bool NANM>
__device__ __forceinline__ uint32_t sd_child(float v, uint32_t thr, uint32_t word, float missing) {
  uint32_t c;
  if (NANM) {
    asm("{\n"
        ".reg .pred pm, pf, pg, pa;\n"
        ".reg .b32 t;\n"
        "setp.nan.f32 pm, %1, %1;\n"
        "and.b32 t, %3, 0x40000000;\n"
        "setp.ne.u32 pf, t, 0;\n"
        "setp.ge.f32 pg, %1, %2;\n"
        "xor.pred pg, pg, pf;\n"
        "setp.lt.s32 pa, %3, 0;\n"
        "and.pred pa, pa, pm;\n"
        "and.pred pg, pg, !pm;\n"
        "or.pred pg, pg, pa;\n"
        "selp.u32 %0, 2, 1, pg;\n"
        "}\n"
        : "=r"(c) : "f"(v), "f"(__uint_as_float(thr)), "r"(word), "f"(missing));
  } else {
    asm("{\n"
        ".reg .pred pm, pf, pg, pa;\n"
        ".reg .b32 t;\n"
        ".reg .f32 d;\n"
        "sub.f32 d, %1, %4;\n"
        "abs.f32 d, d;\n"
        "setp.le.f32 pm, d, 0f358637BD;\n"
        "and.b32 t, %3, 0x40000000;\n"
        "setp.ne.u32 pf, t, 0;\n"
        "setp.ge.f32 pg, %1, %2;\n"
        "xor.pred pg, pg, pf;\n"
        "setp.lt.s32 pa, %3, 0;\n"
        "and.pred pa, pa, pm;\n"
        "and.pred pg, pg, !pm;\n"
        "or.pred pg, pg, pa;\n"
        "selp.u32 %0, 2, 1, pg;\n"
        "}\n"
        : "=r"(c) : "f"(v), "f"(__uint_as_float(thr)), "r"(word), "f"(missing));
  }
  return c;
}

template <int kSdGroup, bool NANM>
__device__ __forceinline__ double sd_walk_row_blocked(const SharedDataParams& p, const uint8_t* top, const uint8_t* my,
                                                      size_t row, int lane, int wt) {
  double acc = 0.0;
  const int gstride = 32 * p.team_warps;
  const uint32_t tpad8 = (uint32_t)p.top_stride * 8u, tpad32 = (uint32_t)p.tpad * 32u;   // (top levels | blocks)
  const float missing = p.missing;
  auto decide = [&](uint32_t val, uint32_t word) -> uint32_t {
    return sd_child<NANM>(*reinterpret_cast<const float*>(my + (word & kSdOffMask)), val, word, missing);
  };
  auto lds_node = [&](uint32_t off) -> uint2 { return *reinterpret_cast<const uint2*>(top + off); };
  const int nslots = p.num_trees;                    // slots 0 .. num_trees-1 hold the trees; the rest is padding
  const uint32_t level_first = (1u << p.top_levels) - 1u;
  for (int s0 = lane + 32 * wt; s0 < nslots; s0 += gstride * kSdGroup) {
    uint32_t n[kSdGroup];      // heap index inside the top, then position inside the current level
    uint32_t slot[kSdGroup];
    uint2 e[kSdGroup];
#pragma unroll
    for (int g = 0; g < kSdGroup; ++g) {
      // past the last tree: repeat its walk (the same addresses as the lane that owns it) and drop the sum
      slot[g] = (uint32_t)min(s0 + gstride * g, nslots - 1);
      n[g] = 0;
      e[g] = lds_node(slot[g] * 8u);
    }
    for (int l = 1; l < p.top_levels; ++l) {          // decisions on levels 0 .. top_levels-2, next node fetched
#pragma unroll
      for (int g = 0; g < kSdGroup; ++g) {
        n[g] = 2 * n[g] + decide(e[g].x, e[g].y);
        e[g] = lds_node(slot[g] * 8u + n[g] * tpad8);
      }
    }
#pragma unroll
    for (int g = 0; g < kSdGroup; ++g)                // the top's last level: position inside level top_levels
      n[g] = 2 * n[g] + decide(e[g].x, e[g].y) - level_first;
    const uint8_t* blocks = reinterpret_cast<const uint8_t*>(p.blocks);
    uint32_t base = 0;
    for (int j = 1; j < p.block_levels; ++j) {        // every trip but the last: two decisions
      uint32_t w[kSdGroup][8];
#pragma unroll
      for (int g = 0; g < kSdGroup; ++g)
        ldg_block(reinterpret_cast<const uint4*>(blocks + (size_t)((base + n[g]) * tpad32 + slot[g] * 32u)), w[g]);
      uint32_t c0[kSdGroup];
#pragma unroll
      for (int g = 0; g < kSdGroup; ++g) c0[g] = decide(w[g][0], w[g][1]);
#pragma unroll
      for (int g = 0; g < kSdGroup; ++g)     // (decide is 1 or 2: child slot + 1)
        n[g] = 4 * n[g] + 2 * c0[g] + decide(c0[g] == 2 ? w[g][4] : w[g][2], c0[g] == 2 ? w[g][5] : w[g][3]) - 3;
      base += 1u << (p.top_levels + 2 * (j - 1));
    }
    float leaf_val[kSdGroup];
    // the last trip: one decision; the children are the bottom-level leaves
    uint32_t w[kSdGroup][8];
#pragma unroll
    for (int g = 0; g < kSdGroup; ++g)
      ldg_block(reinterpret_cast<const uint4*>(blocks + (size_t)((base + n[g]) * tpad32 + slot[g] * 32u)), w[g]);
#pragma unroll
    for (int g = 0; g < kSdGroup; ++g) {
      const uint32_t c0 = decide(w[g][0], w[g][1]);
      leaf_val[g] = __uint_as_float(c0 == 2 ? w[g][4] : w[g][2]);
      n[g] = 2 * n[g] + c0 - 1;
    }
#pragma unroll
    for (int g = 0; g < kSdGroup; ++g) {
      if (s0 + gstride * g >= nslots) continue;
      acc += (double)leaf_val[g];
      if (p.leaf_ids)
        p.leaf_ids[row * (size_t)p.num_trees + p.tree_order[slot[g]]] = p.orig_bottom[(size_t)n[g] * p.tpad + slot[g]];
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  return acc;
}

// Shared-memory map: [teams][row_bufs][row_stride floats] row buffers | [teams][2][4] double partials |
// [teams][2] mbarriers (TMA only) | top levels of the forest (top_off) | the top image's mbarrier.
// TMA = true : rows arrive by bulk copies (cp.async.bulk, row bytes a multiple of 16) completing on
//              the buffer's mbarrier: with two buffers the next row streams in from HBM while the current one is
//              walked; with one it is prefetched into L2 meanwhile and copied after the walk.
// TMA = false: the team copies its row with coalesced loads (any width / alignment).
template <bool TMA, int G, bool BLOCKED>
__global__ void __launch_bounds__(1024, 1) k_shared_data(SharedDataParams p) {
  extern __shared__ __align__(128) uint8_t sd_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int team = warp / p.team_warps, wt = warp % p.team_warps;
  const int team_threads = 32 * p.team_warps;
  const uint32_t row_stride = ((uint32_t)p.cols * 4u + 15u) & ~15u;       // bytes, 16-aligned buffers
  const uint32_t nb = (uint32_t)p.row_bufs;
  float* buf0 = reinterpret_cast<float*>(sd_smem + (size_t)team * nb * row_stride);
  double* partials = reinterpret_cast<double*>(sd_smem + (size_t)p.teams_per_block * nb * row_stride) + team * 8;
  const uint32_t bar0 = smem_u32(sd_smem) + (uint32_t)p.teams_per_block * (nb * row_stride + 64) + team * 16u;
  const size_t team_global = (size_t)blockIdx.x * p.teams_per_block + team;
  const size_t stride = (size_t)gridDim.x * p.teams_per_block;
  const uint32_t row_bytes = (uint32_t)p.cols * 4u;

  auto load_row_tma = [&](size_t row, int b) {       // team warp 0, lane 0
    fence_proxy_async();
    mbar_expect_tx(bar0 + b * 8, row_bytes);
    const uint8_t* src = reinterpret_cast<const uint8_t*>(p.data + row * (size_t)p.cols);
    uint32_t done = 0;
    while (done < row_bytes) {                       // pieces of <= 64 KB
      const uint32_t piece = min(row_bytes - done, 65536u);
      bulk_g2s(smem_u32(buf0) + b * row_stride + done, src + done, piece, bar0 + b * 8);
      done += piece;
    }
  };
  // the team's first row starts to stream in before the CTA copies the forest's top levels
  if (TMA && wt == 0 && lane == 0) {
    mbar_init(bar0, 1); mbar_init(bar0 + 8, 1); fence_mbar_init();
    if (team_global < p.rows) load_row_tma(team_global, 0);
  }
  const uint2* s_top = reinterpret_cast<const uint2*>(sd_smem + p.top_off);
  const uint32_t top_bar = smem_u32(sd_smem) + p.top_off + p.top_bytes;     // (in the 128 spare bytes behind the copy)
  if (BLOCKED) {
    // the packer's image of the top levels ([node][top_stride], a multiple of 16 bytes) by bulk copies, under the
    // teams' first rows
    if (threadIdx.x == 0) {
      mbar_init(top_bar, 1);
      fence_mbar_init();
      mbar_expect_tx(top_bar, p.top_bytes);
      for (uint32_t done = 0; done < p.top_bytes; done += 65536u)
        bulk_g2s(smem_u32(sd_smem) + p.top_off + done, reinterpret_cast<const uint8_t*>(p.top_nodes) + done,
                 min(p.top_bytes - done, 65536u), top_bar);
    }
  } else {
    uint2* dst = reinterpret_cast<uint2*>(sd_smem + p.top_off);
    const int n_top = p.top_levels > 0 ? ((1 << p.top_levels) - 1) * p.tpad : 0;
    for (int i = threadIdx.x; i < n_top; i += blockDim.x) dst[i] = __ldg(p.inter + i);
  }
  __syncthreads();
  if (BLOCKED) mbar_wait(top_bar, 0);
  // (a 16th team on hardware barrier 0 fits shared memory for config 1 and was measured: 0.1453 vs 0.1426 ms with 15)

  // LSU path: short rows wait in registers (one tile ahead) and are stored when their turn comes;
  // rows too long for that are copied synchronously
  constexpr int kStage = 4;
  const int t = wt * 32 + lane;
  const bool staged = p.cols <= kStage * team_threads;
  float stg[kStage];
  auto fetch_row = [&](size_t row) {
    const float* g = p.data + row * (size_t)p.cols;
#pragma unroll
    for (int k = 0; k < kStage; ++k) {
      const int i = t + k * team_threads;
      if (i < p.cols) stg[k] = __ldg(g + i);
    }
  };
  auto commit_row = [&](int b) {
    float* m = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(buf0) + (size_t)b * row_stride);
#pragma unroll
    for (int k = 0; k < kStage; ++k) {
      const int i = t + k * team_threads;
      if (i < p.cols) m[i] = stg[k];
    }
  };
  auto copy_row = [&](size_t row, int b) {
    const float* g = p.data + row * (size_t)p.cols;
    float* m = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(buf0) + (size_t)b * row_stride);
    for (int i = t; i < p.cols; i += team_threads) m[i] = __ldg(g + i);
  };

  // one buffer per team: the row after this one is asked into L2 while this one is walked (no shared memory
  // needed for it), so the team's own bulk copy finds it there instead of waiting for HBM
  auto prefetch_row_l2 = [&](size_t row) {            // one lane
    const uint8_t* src = reinterpret_cast<const uint8_t*>(p.data + row * (size_t)p.cols);
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(row_bytes & ~15u) : "memory");
  };
  const bool l2_ahead = TMA && nb == 1 && p.l2_ahead;
  uint32_t phase = 0;
  int b = 0, pp = 0;
  const bool nan_missing = p.missing != p.missing;
  size_t row = team_global;
  if (!TMA && row < p.rows && staged) {
    fetch_row(row);
  }
  for (; row < p.rows; row += stride) {
    const size_t next = row + stride;
    if (TMA) {
      // buffer b^1 was last read for the previous row; the team barrier at its end released it
      if (nb == 2 && next < p.rows && wt == 0 && lane == 0) load_row_tma(next, b ^ 1);
      if (l2_ahead && next < p.rows && wt == p.team_warps - 1 && lane == 0) prefetch_row_l2(next);
      mbar_wait(bar0 + b * 8, (phase >> b) & 1u);
      phase ^= 1u << b;
    } else {
      if (staged) {
        commit_row(b);
        if (next < p.rows) fetch_row(next);           // lands while this row is walked
      } else {
        copy_row(row, b);
      }
      team_barrier(1 + team, team_threads);           // the team's stores of this row are visible
    }
    const float* my = reinterpret_cast<const float*>(reinterpret_cast<const uint8_t*>(buf0) + (size_t)b * row_stride);
    double acc;
    if (!BLOCKED) acc = sd_walk_row<G>(p, s_top, my, row, lane, wt);
    else if (nan_missing) acc = sd_walk_row_blocked<G, true>(p, sd_smem + p.top_off, reinterpret_cast<const uint8_t*>(my), row, lane, wt);
    else acc = sd_walk_row_blocked<G, false>(p, sd_smem + p.top_off, reinterpret_cast<const uint8_t*>(my), row, lane, wt);
    if (lane == 0) partials[pp * 4 + wt] = acc;    // (two sets: a fast warp's next row cannot overwrite what warp 0 still reads)
    team_barrier(1 + team, team_threads);
    // one buffer: the team barrier above released it; the next row starts to stream in now (issued by the team's
    // LAST warp while its first one writes the score)
    if (TMA && nb == 1 && next < p.rows && wt == p.team_warps - 1 && lane == 0) load_row_tma(next, 0);
    if (wt == 0 && lane == 0) {
      double sum = 0.0;
      for (int w = 0; w < p.team_warps; ++w) sum += partials[pp * 4 + w];   // fixed order: reproducible
      p.preds[row] = apply_transform((float)sum, p.tf);
    }
    b ^= (int)(nb - 1);
    pp ^= 1;
  }
}

int launch_shared_data(tahoe_forest* f, const PredictArgs& a, int* launches) {
  if (f->sparse || !f->d_inter) {
    set_error("SHARED_DATA needs the dense interleaved layout");
    return TAHOE_ERR_NOT_APPLICABLE;
  }
  if (f->sd_teams < 1) {
    set_error("SHARED_DATA: two rows do not fit shared memory");
    return TAHOE_ERR_NOT_APPLICABLE;
  }
  if (a.rows == 0) return TAHOE_OK;
  const size_t row_bytes = (size_t)f->params.num_cols * sizeof(float);
  const size_t row_stride = (row_bytes + 15) & ~(size_t)15;
  SharedDataParams p{};
  p.data = a.data; p.rows = a.rows; p.cols = f->params.num_cols; p.num_trees = f->params.num_trees;
  p.tpad = f->host.tpad; p.depth = f->depth; p.missing = f->params.missing; p.inter = f->d_inter;
  p.orig_ids = f->d_orig_ids; p.tree_order = f->d_tree_order; p.preds = a.out;   // one pass: final scores only
  p.leaf_ids = a.leaf_ids;
  p.tf = Transform{f->params.output, (float)f->params.num_trees, f->params.threshold,
                   f->params.global_bias};
  // the plan (teams, top levels, B32 blocks) was made when the forest was packed: plan_shared_data
  p.team_warps = f->sd_team_warps;
  p.top_levels = f->sd_top;
  p.top_stride = f->sd_top_stride;
  p.teams_per_block = f->sd_teams;
  p.top_nodes = f->d_sd_top_image; p.top_bytes = (uint32_t)(f->host.sd_top_image.size() * sizeof(uint2)); p.blocks = f->d_sd_blocks; p.orig_bottom = f->d_sd_orig_bottom;
  p.block_levels = f->sd_blocked ? (f->depth + 1 - f->sd_top) / 2 : 0;
  p.row_bufs = f->sd_row_bufs;
  p.l2_ahead = f->tune.sd_no_l2_ahead ? 0 : 1;
  const size_t per_team = (size_t)p.row_bufs * row_stride + 64 + 16;    // row buffers, partials, mbarriers
  p.top_off = (uint32_t)(((size_t)p.teams_per_block * per_team + 127) / 128 * 128);
  const size_t smem = f->sd_smem;
  // bulk copies need 16-byte rows at 16-byte addresses; short rows are cheaper through the LSU
  const bool tma = (row_bytes % 16 == 0) && ((reinterpret_cast<uintptr_t>(a.data) & 15) == 0) && row_bytes >= 1024;
  // Small batches are bound by the HOST side of a launch (10 us per call measured against a 5 us kernel,
  // tools/probe_small_batch.py), so what does not change between calls is asked once: the kernel's shared-memory
  // limit per device (raised only when a forest needs more than any before it) and the forest's CTAs per SM.
  auto go = [&](auto kern) -> int {
    // (every instantiation has the same function-pointer TYPE, so this lambda is instantiated once: the limits are
    // keyed by the kernel's address and the device)
    if (smem > 48 * 1024) {
      static std::mutex mu;
      static std::map<std::pair<const void*, int>, int> limit_set;
      std::lock_guard<std::mutex> lock(mu);
      int& have = limit_set[{reinterpret_cast<const void*>(kern), f->device}];
      if ((int)smem > have) {
        TAHOE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        have = (int)smem;
      }
    }
    const int threads = p.teams_per_block * p.team_warps * 32;
    int per_sm = f->sd_per_sm[tma ? 1 : 0];
    if (per_sm == 0) {
      TAHOE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem));
      per_sm = std::max(per_sm, 1);
      f->sd_per_sm[tma ? 1 : 0] = per_sm;
    }
    const size_t want = (a.rows + p.teams_per_block - 1) / p.teams_per_block;
    const unsigned grid = (unsigned)std::min<size_t>(want, (size_t)f->sm_count * per_sm);
    kern<<<grid, threads, smem, a.stream>>>(p);
    TAHOE_CUDA(cudaGetLastError());
    ++*launches;
    return TAHOE_OK;
  };
  const int G = f->sd_group;
  if (f->sd_blocked) {
    if (G == 4) return tma ? go(k_shared_data<true, 4, true>) : go(k_shared_data<false, 4, true>);
    if (G == 2) return tma ? go(k_shared_data<true, 2, true>) : go(k_shared_data<false, 2, true>);
    return tma ? go(k_shared_data<true, 1, true>) : go(k_shared_data<false, 1, true>);
  }
  if (G == 4) return tma ? go(k_shared_data<true, 4, false>) : go(k_shared_data<false, 4, false>);
  if (G == 2) return tma ? go(k_shared_data<true, 2, false>) : go(k_shared_data<false, 2, false>);
  return tma ? go(k_shared_data<true, 1, false>) : go(k_shared_data<false, 1, false>);
}

}  // namespace tahoe
