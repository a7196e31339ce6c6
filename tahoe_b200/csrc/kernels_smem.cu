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
#ifndef TAHOE_GROUP
#define TAHOE_GROUP 4
#endif
constexpr int kGroup = TAHOE_GROUP;  // trees traversed together by one thread (ILP)

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
  int rank_shift;
  uint32_t miss_code;
  uint32_t* qtiles;        // [tiles][cols][32]
  uint32_t* tile_flags;    // [tiles]
  size_t num_tiles;
};

constexpr int kQuantThreads = 256;

// Persistent blocks; each keeps the bucket LUT of every column in shared memory.  A warp owns a
// 32-row tile: its input is one contiguous run of 32*cols floats (coalesced reads), its output the
// column-major tile [col][32].  rank(v) = #{thr <= v}: one LUT lookup picks v's bucket, a short
// binary search over that bucket's thresholds (one or two cache lines) finishes it.
__global__ void __launch_bounds__(kQuantThreads) k_quantize(QuantParams p) {
  extern __shared__ __align__(16) uint8_t qsm[];
  const int B1 = p.buckets + 1;
  int32_t* s_lut = reinterpret_cast<int32_t*>(qsm);                    // [cols][B+1]
  float* s_min = reinterpret_cast<float*>(s_lut + (size_t)p.cols * B1);  // [cols]
  float* s_scale = s_min + p.cols;                                     // [cols]
  int32_t* s_off = reinterpret_cast<int32_t*>(s_scale + p.cols);       // [cols+1]
  for (int i = threadIdx.x; i < p.cols * B1; i += kQuantThreads) s_lut[i] = p.lut[i];
  for (int i = threadIdx.x; i < p.cols; i += kQuantThreads) { s_min[i] = p.lut_min[i]; s_scale[i] = p.lut_scale[i]; }
  for (int i = threadIdx.x; i <= p.cols; i += kQuantThreads) s_off[i] = p.offsets[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const size_t warp = ((size_t)blockIdx.x * kQuantThreads + threadIdx.x) >> 5;
  const size_t nwarps = ((size_t)gridDim.x * kQuantThreads) >> 5;
  const int step_r = 32 / p.cols, step_c = 32 % p.cols;   // advance of (row, col) per 32 words
  const float fB = (float)p.buckets;
  for (size_t tile = warp; tile < p.num_tiles; tile += nwarps) {
    const size_t row0 = tile * 32;
    const size_t nrows = min((size_t)32, p.rows - row0);
    const int total = (int)nrows * p.cols;
    const float* src = p.data + row0 * (size_t)p.cols;
    uint32_t* dst = p.qtiles + tile * (size_t)p.cols * 32;
    bool any_missing = false;
    int r = lane / p.cols, c = lane - r * p.cols;
    for (int w = lane; w < 32 * p.cols; w += 32) {
      uint32_t code = 0;
      if (w < total) {
        const float v = __ldg(src + w);
        if (fabsf(v - p.missing) <= 1.0e-6f) {      // BaseTahoeTest.h:451-452
          code = p.miss_code;
          any_missing = true;
        } else if (v == v) {                         // NaN: val >= thr is never true -> rank 0
          const float x = (v - s_min[c]) * s_scale[c];
          int b = x >= fB ? p.buckets - 1 : (int)x;  // float->int truncation is monotone
          b = max(b, 0);
          const int32_t* lut = s_lut + c * B1;
          const float* tab = p.table + s_off[c];
          int lo = lut[b], hi = lut[b + 1];
          while (lo < hi) {                          // rank = #{thr <= v}
            const int mid = (lo + hi) >> 1;
            if (__ldg(tab + mid) <= v) lo = mid + 1; else hi = mid;
          }
          code = (uint32_t)lo << p.rank_shift;
        }
      }
      dst[c * 32 + r] = code;
      r += step_r; c += step_c;
      if (c >= p.cols) { c -= p.cols; ++r; }
    }
    any_missing = __any_sync(0xffffffffu, any_missing);
    if (lane == 0) p.tile_flags[tile] = any_missing ? 1u : 0u;
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
  int num_trees, trees_per_chunk, num_chunks;
  uint32_t col_mask, miss_code;
  float* preds;
  int32_t* leaf_ids;
  const int32_t* leaf_origin;  // [T][2^d]
  Transform tf;
};

constexpr int kSlots = kSmemWarps * kSmemTileBufs;     // tile buffers per CTA (32)
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

template <int DEPTH, bool LEAF_IDS, bool WIDE>
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
    const int ntrees = min(p.trees_per_chunk, p.num_trees - tree0);
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
    for (; tile < p.num_tiles; tile += warp_stride) {
      // prefetch the next tile into the other buffer (all lanes finished reading it)
      const size_t next = tile + warp_stride;
      __syncwarp();
      if (next < p.num_tiles) load_tile(next, buf ^ 1);
      const bool has_missing = __ldg(p.tile_flags + tile) != 0;
      mbar_wait(bar_tile0 + buf * 8, (tile_phase >> buf) & 1u);
      tile_phase ^= 1u << buf;

      const size_t row = tile * 32 + lane;
      const bool live = row < p.rows;
      float acc = (first || !live) ? 0.0f : p.preds[row];
      const uint32_t slot_lane = (uint32_t)(warp * kSmemTileBufs + buf) * 128u + lane * 4u;

      int t = 0;
      for (; t + kGroup <= ntrees; t += kGroup) {
        uint32_t a[kGroup];
        const uint32_t tree_abs = s_base + forest_off + (uint32_t)t * W * 4u;
        if (has_missing) traverse_group<DEPTH, kGroup, true, WIDE>(tree_abs, s_base, slot_lane, p.col_mask, p.miss_code, a);
        else traverse_group<DEPTH, kGroup, false, WIDE>(tree_abs, s_base, slot_lane, p.col_mask, p.miss_code, a);
#pragma unroll
        for (int g = 0; g < kGroup; ++g) {
          acc += __uint_as_float(lds_u32(a[g]));
          if (LEAF_IDS && live)
            p.leaf_ids[row * (size_t)p.num_trees + tree0 + t + g] =
                __ldg(p.leaf_origin + (size_t)(tree0 + t + g) * L + ((a[g] - tree_abs - g * W * 4u) / 4 - L));
        }
      }
      for (; t < ntrees; ++t) {
        uint32_t a[1];
        const uint32_t tree_abs = s_base + forest_off + (uint32_t)t * W * 4u;
        if (has_missing) traverse_group<DEPTH, 1, true, WIDE>(tree_abs, s_base, slot_lane, p.col_mask, p.miss_code, a);
        else traverse_group<DEPTH, 1, false, WIDE>(tree_abs, s_base, slot_lane, p.col_mask, p.miss_code, a);
        acc += __uint_as_float(lds_u32(a[0]));
        if (LEAF_IDS && live)
          p.leaf_ids[row * (size_t)p.num_trees + tree0 + t] =
              __ldg(p.leaf_origin + (size_t)(tree0 + t) * L + ((a[0] - tree_abs) / 4 - L));
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
  if (p.leaf_ids) return wide ? go(k_smem_forest<DEPTH, true, true>) : go(k_smem_forest<DEPTH, true, false>);
  return wide ? go(k_smem_forest<DEPTH, false, true>) : go(k_smem_forest<DEPTH, false, false>);
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
  q.rank_shift = f->rank_shift; q.miss_code = f->miss_code; q.qtiles = a.ws->qtiles;
  q.tile_flags = a.ws->tile_flags; q.num_tiles = tiles;
  const size_t qsmem = ((size_t)q.cols * (q.buckets + 1) + 3 * (size_t)q.cols + 1) * 4;
  if (qsmem > 48 * 1024)
    TAHOE_CUDA(cudaFuncSetAttribute(k_quantize, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qsmem));
  const size_t qwarps_per_block = kQuantThreads / 32;
  const int qblocks_per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, ((size_t)f->smem_optin - 1024) / (qsmem + 1024)));
  unsigned qgrid = (unsigned)std::min<size_t>((tiles + qwarps_per_block - 1) / qwarps_per_block,
                                              (size_t)f->sm_count * qblocks_per_sm);
  cudaEvent_t* ev = nullptr;
  if (f->profile && a.ws == &f->ws[0] && f->prof_count < kProfRing) {
    ev = &f->prof_events[(size_t)f->prof_count * 3];
    ++f->prof_count;
    TAHOE_CUDA(cudaEventRecord(ev[0], a.stream));
  }
  k_quantize<<<qgrid, kQuantThreads, qsmem, a.stream>>>(q);
  if (ev) TAHOE_CUDA(cudaEventRecord(ev[1], a.stream));
  TAHOE_CUDA(cudaGetLastError());
  ++*launches;

  SmemParams p{};
  p.qtiles = a.ws->qtiles; p.tile_flags = a.ws->tile_flags; p.num_tiles = tiles; p.rows = a.rows;
  p.cols = f->params.num_cols; p.qwords = f->d_qwords; p.num_trees = f->params.num_trees;
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
