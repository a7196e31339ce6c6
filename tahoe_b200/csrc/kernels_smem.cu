// kernels_smem.cu -- SHARED_FOREST / SPLIT_SHARED_FOREST (reference K4/K5,
// Struct.h:1245-1409 and 1414-1606), rebuilt for sm_100a:
//
//   k_quantize     one pass over the float rows: every feature becomes its RANK among
//                  the forest's distinct thresholds on that column (exact: v >= thr_k
//                  <=> rank(v) > k), stored per 32-row warp tile, column-major
//                  [tile][col][32 lanes] so one bulk copy brings a tile in and lane i
//                  only ever touches bank i.
//   k_smem_forest  persistent CTAs (one per SM, 227 KB shared memory).  The forest is
//                  cut into chunks of whole trees; for each chunk the CTA bulk-copies it
//                  into shared memory ONCE (the reference re-copies the whole forest in
//                  every block) and streams all of its row tiles through it.
//                  Tiles arrive by cp.async.bulk + mbarrier, double-buffered per warp.
//                  thread = sample, 4 trees in flight per thread; a node visit is one LDS
//                  (node word), one LDS (feature rank) and one unsigned compare.
//                  Partial sums are carried in preds[] between chunks by the same thread,
//                  in tree order, so the float sum is the CPU's sum bit for bit -- no
//                  [rows x blocks] temp and no segmented reduce.
#include "device_utils.cuh"
#include "forest.h"

namespace tahoe {

constexpr int kSmemThreads = kSmemWarps * 32;

// ---------------------------------------------------------------------------------------------
struct QuantParams {
  const float* data;
  size_t rows;
  int cols;
  float missing;
  const float* table;      // sorted distinct thresholds, per column
  const int32_t* offsets;  // [cols+1]
  int buckets;             // B
  const float* lut_min;    // [cols]
  const float* lut_scale;  // [cols]
  const int32_t* lut;      // [cols][B+1]
  int stride_log2;         // shared memory keeps every 2^stride_log2-th threshold of the column
  int rank_shift;
  uint32_t miss_code;
  uint32_t* qtiles;        // [tiles][cols][32]
  uint32_t* tile_flags;    // [tiles], zeroed before the launch
  size_t num_tiles;
  int splits;              // CTAs per column
};

constexpr int kQuantThreads = 1024;
constexpr int kQuantUnroll = 4;   // tiles in flight per warp (independent strided loads)

// One CTA = one COLUMN of the sample matrix over a range of 32-row tiles.  The column's sorted
// threshold table (112 KB for the SUSY-shaped forest) and its bucket index sit in shared memory, so
// rank(v) = #{thr <= v} is one index lookup plus a two-or-three-probe search that never leaves the
// SM.  The CTAs of all columns walk the same rows at the same pace: a 32-byte sector of the
// row-major input is fetched from HBM once and served to the other columns' CTAs by L2.  Output is
// the column-major tile [tile][col][32 lanes]: one coalesced 128-byte store per warp.
// Columns longer than shared memory keep every 2^k-th threshold there and finish with <= 2^k - 1
// probes of one or two sectors of the full table.
__global__ void __launch_bounds__(kQuantThreads, 1) k_quantize(QuantParams p) {
  extern __shared__ __align__(16) uint8_t qsm[];
  const int c = blockIdx.x % p.cols;
  const int split = blockIdx.x / p.cols;
  const int B = p.buckets;
  const int S = 1 << p.stride_log2;
  const int off = p.offsets[c];
  const int n = p.offsets[c + 1] - off;
  const int ns = n >> p.stride_log2;                 // whole blocks of S thresholds
  int32_t* s_lut = reinterpret_cast<int32_t*>(qsm);  // [B+1], in units of blocks, clamped to ns
  float* s_tab = reinterpret_cast<float*>(s_lut + (B + 1));   // [ns] last threshold of each block
  const float* tab = p.table + off;
  for (int i = threadIdx.x; i <= B; i += kQuantThreads)
    s_lut[i] = min(ns, p.lut[(size_t)c * (B + 1) + i] >> p.stride_log2);
  for (int i = threadIdx.x; i < ns; i += kQuantThreads) s_tab[i] = __ldg(tab + (size_t)i * S + (S - 1));
  __syncthreads();
  const float mn = p.lut_min[c], scale = p.lut_scale[c], fB = (float)B;

  auto rank_code = [&](float v, bool& any_missing) -> uint32_t {
    if (fabsf(v - p.missing) <= 1.0e-6f) {      // BaseTahoeTest.h:451-452
      any_missing = true;
      return p.miss_code;
    }
    if (!(v == v)) return 0u;                    // NaN: val >= thr is never true -> rank 0
    const float x = (v - mn) * scale;
    int b = x >= fB ? B - 1 : (int)x;            // float->int truncation is monotone
    b = max(b, 0);
    int lo = s_lut[b], hi = s_lut[b + 1];
    while (lo < hi) {                            // first block whose last threshold is > v
      const int mid = (lo + hi) >> 1;
      if (s_tab[mid] <= v) lo = mid + 1; else hi = mid;
    }
    int r = lo << p.stride_log2;
    if (S > 1) {                                 // count inside that block (sorted: stop at the first > v)
      const int end = min(n, r + S);
      while (r < end && __ldg(tab + r) <= v) ++r;
    }
    return (uint32_t)r << p.rank_shift;
  };

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int kWarps = kQuantThreads / 32;
  const size_t per = (p.num_tiles + p.splits - 1) / p.splits;
  const size_t t0 = (size_t)split * per, t1 = min(p.num_tiles, t0 + per);
  for (size_t base = t0 + warp; base < t1; base += (size_t)kWarps * kQuantUnroll) {
    float v[kQuantUnroll];
    bool live[kQuantUnroll];
#pragma unroll
    for (int u = 0; u < kQuantUnroll; ++u) {
      const size_t tile = base + (size_t)u * kWarps;
      const size_t row = tile * 32 + lane;
      live[u] = tile < t1 && row < p.rows;
      v[u] = live[u] ? __ldg(p.data + row * (size_t)p.cols + c) : 0.0f;
    }
#pragma unroll
    for (int u = 0; u < kQuantUnroll; ++u) {
      const size_t tile = base + (size_t)u * kWarps;
      if (tile >= t1) break;                     // warp-uniform
      bool miss = false;
      const uint32_t code = live[u] ? rank_code(v[u], miss) : 0u;
      p.qtiles[(tile * p.cols + c) * 32 + lane] = code;
      if (__any_sync(0xffffffffu, miss) && lane == 0) atomicOr(p.tile_flags + tile, 1u);
    }
  }
}

// ---------------------------------------------------------------------------------------------
struct SmemParams {
  const uint32_t* qtiles;
  const uint32_t* tile_flags;
  size_t num_tiles;
  size_t rows;
  int cols;
  const uint32_t* qwords;   // [T][2^(d+1)]
  int num_trees;            // real trees (leaf ids are written for these only)
  int padded_trees;         // trees in qwords: whole traversal groups, the extra ones all zero
  int trees_per_chunk, num_chunks;
  uint32_t col_mask, miss_code;
  float* preds;
  int32_t* leaf_ids;
  const int32_t* leaf_origin;  // [T][2^d]
  Transform tf;
};

constexpr uint32_t kColStride = 128u * kSlots;          // bytes between column rows in the tile region
static_assert(kColStride == (1u << kQuantColShiftWide), "wide column field must equal the region stride");

// Shared-memory map of k_smem_forest (byte offsets from smem[]):
//   [0, cols*kColStride)           tile region: word (col, slot, lane) at col*4096 + slot*128 + lane*4
//                                  -- all 32 warp buffers interleaved, so the region base is one
//                                  CTA-uniform constant and a lane's address is a pure bit-field OR
//   [.., + trees_per_chunk*W*4)    the resident chunk of rank-quantised trees
//   [.., + (1 + kSlots)*8)         mbarriers: chunk, then one per tile buffer

// One group of G trees for this lane's sample, on absolute shared-memory addresses.
//   tree_abs   shared address of word 0 of the group's first tree (warp-uniform)
//   s_base     shared address of smem[0] = base of the tile region (CTA-uniform)
//   slot_lane  slot*128 + lane*4 : this lane's bank inside this warp's current tile buffer
// a[g] is the shared ADDRESS of the current node, so one visit is six instructions:
//   LDS  node  [a]
//   LOP3 (node & col_mask) | slot_lane           (WIDE; the compact form adds one shift)
//   LDS  rank  [that + s_base]                   (uniform-register base, folded into the LDS)
//   ISETP rank > node ; SEL k = right ? 4-B : -B ; IADD3 a = a + a + k      (B = tree address)
// i.e. a' - B = 2(a - B) + 4*right, the heap step on byte offsets, with no separate address add.
// Loads are asm so the address arithmetic stays exactly this; ptxas still interleaves the G chains.
template <int DEPTH, int G, bool MISSING, bool WIDE>
__device__ __forceinline__ void traverse_group(uint32_t tree_abs, uint32_t s_base, uint32_t slot_lane,
                                               uint32_t col_mask, uint32_t miss_code, uint32_t (&a)[G]) {
  constexpr uint32_t W4 = 4u << (DEPTH + 1);  // bytes per tree
  uint32_t k0[G];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    k0[g] = 0u - (tree_abs + g * W4);
    a[g] = tree_abs + g * W4 + 4;
  }
#pragma unroll
  for (int l = 0; l < DEPTH; ++l) {
    uint32_t N[G], F[G];
#pragma unroll
    for (int g = 0; g < G; ++g) N[g] = lds_u32(a[g]);
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const uint32_t col_off = WIDE ? (N[g] & col_mask)
                                    : ((N[g] & col_mask) << (kQuantColShiftWide - kQuantColShift));
      F[g] = lds_u32(s_base + (col_off | slot_lane));
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      bool c = F[g] > N[g];
      if (MISSING) {
        if (F[g] == miss_code) c = !(N[g] & 1u);  // sentinel: !def_left (BaseTahoeTest.h:452)
      }
      a[g] = a[g] + a[g] + (c ? k0[g] + 4u : k0[g]);
    }
  }
}

template <int DEPTH, int G, bool LEAF_IDS, bool WIDE>
__global__ void __launch_bounds__(kSmemThreads, 1) k_smem_forest(SmemParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr uint32_t W = 1u << (DEPTH + 1);      // words per tree
  constexpr uint32_t L = 1u << DEPTH;            // bottom slots per tree
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t tile_bytes = (uint32_t)p.cols * 128u;
  const uint32_t forest_off = (uint32_t)p.cols * kColStride;
  const uint32_t bars_off = forest_off + (uint32_t)p.trees_per_chunk * W * 4u;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t bar_chunk = s_base + bars_off;
  const uint32_t bar_tile0 = bar_chunk + 8 + warp * (8 * kSmemTileBufs);

  if (threadIdx.x == 0) {
    mbar_init(bar_chunk, 1);
    for (int i = 0; i < kSlots; ++i) mbar_init(bar_chunk + 8 + i * 8, 1);
    fence_mbar_init();
  }
  __syncthreads();

  const size_t warp_global = (size_t)blockIdx.x * kSmemWarps + warp;
  const size_t warp_stride = (size_t)gridDim.x * kSmemWarps;
  uint32_t tile_phase = 0;   // bit b = parity to wait for on this warp's buffer b
  uint32_t chunk_phase = 0;

  // bring tile `tile` into this warp's buffer `b`: one 128-byte bulk copy per column row
  auto load_tile = [&](size_t tile, int b) {
    const uint32_t bar = bar_tile0 + b * 8;
    if (lane == 0) mbar_expect_tx(bar, tile_bytes);
    __syncwarp();
    const uint32_t slot_off = (uint32_t)(warp * kSmemTileBufs + b) * 128u;
    const uint32_t* src = p.qtiles + tile * (size_t)p.cols * 32;
    for (int c = lane; c < p.cols; c += 32) {
      fence_proxy_async();
      bulk_g2s(s_base + c * kColStride + slot_off, src + c * 32, 128u, bar);
    }
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
    if (tile < p.num_tiles) load_tile(tile, 0);
    mbar_wait(bar_chunk, chunk_phase);
    chunk_phase ^= 1;

    const bool first = (chunk == 0), last = (chunk == p.num_chunks - 1);
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
      // prefetch the next tile into the other buffer (all lanes finished reading it)
      const size_t next = tile + warp_stride;
      __syncwarp();
      const bool has_missing = flag_next != 0;
      float acc = acc_next;
      if (next < p.num_tiles) {
        load_tile(next, buf ^ 1);
        flag_next = __ldg(p.tile_flags + next);
        const size_t rn = next * 32 + lane;
        acc_next = (!first && rn < p.rows) ? p.preds[rn] : 0.0f;
      }
      mbar_wait(bar_tile0 + buf * 8, (tile_phase >> buf) & 1u);
      tile_phase ^= 1u << buf;

      const size_t row = tile * 32 + lane;
      const bool live = row < p.rows;
      const uint32_t slot_lane = (uint32_t)(warp * kSmemTileBufs + buf) * 128u + lane * 4u;

      for (int t = 0; t < ntrees; t += G) {
        uint32_t a[G];
        const uint32_t tree_abs = s_base + forest_off + (uint32_t)t * W * 4u;
        if (has_missing) traverse_group<DEPTH, G, true, WIDE>(tree_abs, s_base, slot_lane, p.col_mask, p.miss_code, a);
        else traverse_group<DEPTH, G, false, WIDE>(tree_abs, s_base, slot_lane, p.col_mask, p.miss_code, a);
#pragma unroll
        for (int g = 0; g < G; ++g) {
          acc += __uint_as_float(lds_u32(a[g]));   // a padding tree adds +0.0f: the sum keeps its bits
          if (LEAF_IDS && live && tree0 + t + g < p.num_trees)
            p.leaf_ids[row * (size_t)p.num_trees + tree0 + t + g] =
                __ldg(p.leaf_origin + (size_t)(tree0 + t + g) * L + ((a[g] - tree_abs - g * W * 4u) / 4 - L));
        }
      }
      if (live) p.preds[row] = last ? apply_transform(acc, p.tf) : acc;
      buf ^= 1;
    }
  }
}

// ---------------------------------------------------------------------------------------------
int ensure_workspace(tahoe_forest* f, Workspace* ws, size_t rows) {
  const size_t tiles = (rows + 31) / 32;
  if (tiles <= ws->cap_tiles) return TAHOE_OK;
  if (ws->qtiles) cudaFree(ws->qtiles);
  if (ws->tile_flags) cudaFree(ws->tile_flags);
  ws->qtiles = nullptr; ws->tile_flags = nullptr; ws->cap_tiles = 0;
  TAHOE_CUDA(cudaMalloc((void**)&ws->qtiles, tiles * (size_t)f->params.num_cols * 128));
  TAHOE_CUDA(cudaMalloc((void**)&ws->tile_flags, tiles * sizeof(uint32_t)));
  ws->cap_tiles = tiles;
  return TAHOE_OK;
}

template <int DEPTH>
static int launch_depth(tahoe_forest* f, const SmemParams& p, unsigned grid, cudaStream_t stream) {
  auto go = [&](auto kern) -> int {
    TAHOE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f->smem_bytes));
    kern<<<grid, kSmemThreads, f->smem_bytes, stream>>>(p);
    TAHOE_CUDA(cudaGetLastError());
    return TAHOE_OK;
  };
  const bool wide = f->col_shift == kQuantColShiftWide;
  // the leaf-id instantiation is the checking path: one tree at a time keeps the build small
  if (p.leaf_ids) return wide ? go(k_smem_forest<DEPTH, 1, true, true>) : go(k_smem_forest<DEPTH, 1, true, false>);
  if (f->group == kGroup) return wide ? go(k_smem_forest<DEPTH, kGroup, false, true>) : go(k_smem_forest<DEPTH, kGroup, false, false>);
  if (f->group >= 2) return wide ? go(k_smem_forest<DEPTH, 2, false, true>) : go(k_smem_forest<DEPTH, 2, false, false>);
  return wide ? go(k_smem_forest<DEPTH, 1, false, true>) : go(k_smem_forest<DEPTH, 1, false, false>);
}

int launch_smem_forest(tahoe_forest* f, const PredictArgs& a, int* launches) {
  if (!f->quant || f->trees_per_chunk < 1) {
    set_error("shared-memory forest strategies need a dense forest of depth 1..14, <= 64 columns, "
              "whose rank+column code fits 32 bits and whose trees fit shared memory");
    return TAHOE_ERR_NOT_APPLICABLE;
  }
  if (a.rows == 0) return TAHOE_OK;
  int rc = ensure_workspace(f, a.ws, a.rows);
  if (rc) return rc;
  const size_t tiles = (a.rows + 31) / 32;

  QuantParams q{};
  q.data = a.data; q.rows = a.rows; q.cols = f->params.num_cols; q.missing = f->params.missing;
  q.table = f->d_qtable; q.offsets = f->d_qoffsets; q.buckets = f->host.qbuckets;
  q.lut_min = f->d_qlut_min; q.lut_scale = f->d_qlut_scale; q.lut = f->d_qlut;
  q.stride_log2 = ilog2_c(f->quant_stride);
  q.rank_shift = f->rank_shift; q.miss_code = f->miss_code; q.qtiles = a.ws->qtiles;
  q.tile_flags = a.ws->tile_flags; q.num_tiles = tiles;
  // CTAs = columns x splits; every split is a contiguous range of tiles, at least one pass of the
  // CTA's warps long, and the grid fills the SMs once (one CTA per SM: its table owns shared memory)
  const size_t tiles_per_pass = (size_t)(kQuantThreads / 32) * kQuantUnroll;
  q.splits = (int)std::max<size_t>(1, std::min<size_t>((tiles + tiles_per_pass - 1) / tiles_per_pass,
                                                        (size_t)std::max(1, f->sm_count / q.cols)));
  const unsigned qgrid = (unsigned)(q.cols * q.splits);
  if (f->quant_smem > 48 * 1024)
    TAHOE_CUDA(cudaFuncSetAttribute(k_quantize, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f->quant_smem));
  cudaEvent_t* ev = nullptr;
  if (f->profile && a.ws == &f->ws[0] && f->prof_count < kProfRing) {
    ev = &f->prof_events[(size_t)f->prof_count * 3];
    ++f->prof_count;
    TAHOE_CUDA(cudaEventRecord(ev[0], a.stream));
  }
  TAHOE_CUDA(cudaMemsetAsync(a.ws->tile_flags, 0, tiles * sizeof(uint32_t), a.stream));
  k_quantize<<<qgrid, kQuantThreads, f->quant_smem, a.stream>>>(q);
  if (ev) TAHOE_CUDA(cudaEventRecord(ev[1], a.stream));
  TAHOE_CUDA(cudaGetLastError());
  ++*launches;

  SmemParams p{};
  p.qtiles = a.ws->qtiles; p.tile_flags = a.ws->tile_flags; p.num_tiles = tiles; p.rows = a.rows;
  p.cols = f->params.num_cols; p.qwords = f->d_qwords; p.num_trees = f->params.num_trees;
  p.padded_trees = f->qwords_trees;
  p.trees_per_chunk = f->trees_per_chunk; p.num_chunks = f->num_chunks;
  p.col_mask = f->col_mask; p.miss_code = f->miss_code; p.preds = a.preds; p.leaf_ids = a.leaf_ids;
  p.leaf_origin = f->d_leaf_origin;
  p.tf = Transform{f->params.output, (float)f->params.num_trees, f->params.threshold,
                   f->params.global_bias};
  const unsigned grid = (unsigned)std::min<size_t>((size_t)f->sm_count,
                                                   (tiles + kSmemWarps - 1) / kSmemWarps);
  switch (f->depth) {
#define TAHOE_DEPTH_CASE(D) case D: rc = launch_depth<D>(f, p, grid, a.stream); break;
    TAHOE_DEPTH_CASE(1) TAHOE_DEPTH_CASE(2) TAHOE_DEPTH_CASE(3) TAHOE_DEPTH_CASE(4)
    TAHOE_DEPTH_CASE(5) TAHOE_DEPTH_CASE(6) TAHOE_DEPTH_CASE(7) TAHOE_DEPTH_CASE(8)
    TAHOE_DEPTH_CASE(9) TAHOE_DEPTH_CASE(10) TAHOE_DEPTH_CASE(11) TAHOE_DEPTH_CASE(12)
    TAHOE_DEPTH_CASE(13) TAHOE_DEPTH_CASE(14)
#undef TAHOE_DEPTH_CASE
    default: set_error("unsupported depth"); return TAHOE_ERR_NOT_APPLICABLE;
  }
  if (rc) return rc;
  if (ev) TAHOE_CUDA(cudaEventRecord(ev[2], a.stream));
  ++*launches;
  return TAHOE_OK;
}

}  // namespace tahoe
