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
constexpr int kGroup = 4;  // trees traversed together by one thread (ILP)

// ---------------------------------------------------------------------------------------------
struct QuantParams {
  const float* data;
  size_t rows;
  int cols;
  float missing;
  const float* table;     // sorted distinct thresholds, per column
  const int32_t* offsets; // [cols+1]
  int rank_shift;
  uint32_t miss_code;
  uint32_t* qtiles;       // [tiles][cols][32]
  uint32_t* tile_flags;   // [tiles]
  size_t num_tiles;
};

__global__ void __launch_bounds__(256) k_quantize(QuantParams p) {
  extern __shared__ int32_t s_off[];  // [cols+1]
  for (int i = threadIdx.x; i <= p.cols; i += blockDim.x) s_off[i] = p.offsets[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const size_t nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
  for (size_t tile = warp; tile < p.num_tiles; tile += nwarps) {
    const size_t row0 = tile * 32;
    const size_t nrows = min((size_t)32, p.rows - row0);
    const int total = (int)nrows * p.cols;
    const float* src = p.data + row0 * (size_t)p.cols;
    uint32_t* dst = p.qtiles + tile * (size_t)p.cols * 32;
    bool any_missing = false;
    for (int w = lane; w < 32 * p.cols; w += 32) {
      const int r = w / p.cols, c = w - r * p.cols;
      uint32_t code = 0;
      if (w < total) {
        const float v = __ldg(src + w);
        if (fabsf(v - p.missing) <= 1.0e-6f) {      // BaseTahoeTest.h:451-452
          code = p.miss_code;
          any_missing = true;
        } else {
          const float* tab = p.table + s_off[c];
          int lo = 0, hi = s_off[c + 1] - s_off[c];
          while (lo < hi) {                          // rank = #{thr <= v}; NaN compares false -> 0
            const int mid = (lo + hi) >> 1;
            if (__ldg(tab + mid) <= v) lo = mid + 1; else hi = mid;
          }
          code = (uint32_t)lo << p.rank_shift;
        }
      }
      dst[c * 32 + r] = code;
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

// One group of G trees for this lane's sample.  tree: shared pointer to word 0 of the
// first tree; tile_lane: shared pointer to (this warp's tile) + lane*4 bytes.
// Plain shared-memory loads (not asm) so that ptxas can interleave the G chains.
template <int DEPTH, int G, bool MISSING>
__device__ __forceinline__ void traverse_group(const uint32_t* tree, const uint8_t* tile_lane,
                                               uint32_t col_mask, uint32_t miss_code,
                                               uint32_t (&n)[G]) {
  constexpr uint32_t W = 1u << (DEPTH + 1);  // words per tree
#pragma unroll
  for (int g = 0; g < G; ++g) n[g] = 1;
#pragma unroll
  for (int l = 0; l < DEPTH; ++l) {
    uint32_t N[G], F[G];
#pragma unroll
    for (int g = 0; g < G; ++g) N[g] = tree[g * W + n[g]];
#pragma unroll
    for (int g = 0; g < G; ++g) F[g] = *(const uint32_t*)(tile_lane + (N[g] & col_mask));
#pragma unroll
    for (int g = 0; g < G; ++g) {
      bool c = F[g] > N[g];
      if (MISSING) {
        if (F[g] == miss_code) c = !(N[g] & 1u);  // sentinel: !def_left (BaseTahoeTest.h:452)
      }
      n[g] = 2 * n[g] + (c ? 1u : 0u);
    }
  }
}

template <int DEPTH, bool LEAF_IDS>
__global__ void __launch_bounds__(kSmemThreads, 1) k_smem_forest(SmemParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr uint32_t W = 1u << (DEPTH + 1);      // words per tree
  constexpr uint32_t L = 1u << DEPTH;            // bottom slots per tree
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t tile_bytes = (uint32_t)p.cols * 128u;
  // layout: [warp tiles: kSmemWarps * 2 * tile_bytes][forest chunk][barriers]
  const uint32_t s_base = smem_u32(smem);
  const uint32_t s_tiles = s_base;
  const uint32_t s_forest = s_tiles + kSmemWarps * kSmemTileBufs * tile_bytes;
  const uint32_t s_bars = s_forest + (uint32_t)p.trees_per_chunk * W * 4u;
  const uint32_t bar_chunk = s_bars;                            // 1 barrier
  const uint32_t bar_tile = s_bars + 8 + warp * (8 * kSmemTileBufs);  // 2 per warp

  if (threadIdx.x == 0) {
    mbar_init(bar_chunk, 1);
    for (int w = 0; w < kSmemWarps; ++w)
      for (int b = 0; b < kSmemTileBufs; ++b) mbar_init(s_bars + 8 + w * (8 * kSmemTileBufs) + b * 8, 1);
    fence_mbar_init();
  }
  __syncthreads();

  const size_t warp_global = (size_t)blockIdx.x * kSmemWarps + warp;
  const size_t warp_stride = (size_t)gridDim.x * kSmemWarps;
  const uint32_t my_tiles = s_tiles + warp * (kSmemTileBufs * tile_bytes);
  uint32_t tile_phase = 0;   // bit b = parity to wait for on buffer b
  uint32_t chunk_phase = 0;

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
        bulk_g2s(s_forest + done, src + done, piece, bar_chunk);
        done += piece;
      }
    }
    // first tile of this pass can fly while the chunk lands
    size_t tile = warp_global;
    if (lane == 0 && tile < p.num_tiles) {
      fence_proxy_async();
      mbar_expect_tx(bar_tile, tile_bytes);
      bulk_g2s(my_tiles, p.qtiles + tile * (size_t)p.cols * 32, tile_bytes, bar_tile);
    }
    mbar_wait(bar_chunk, chunk_phase);
    chunk_phase ^= 1;

    const bool first = (chunk == 0), last = (chunk == p.num_chunks - 1);
    int buf = 0;
    for (; tile < p.num_tiles; tile += warp_stride) {
      // prefetch the next tile into the other buffer (all lanes finished reading it)
      const size_t next = tile + warp_stride;
      __syncwarp();
      if (lane == 0 && next < p.num_tiles) {
        fence_proxy_async();
        mbar_expect_tx(bar_tile + (buf ^ 1) * 8, tile_bytes);
        bulk_g2s(my_tiles + (buf ^ 1) * tile_bytes, p.qtiles + next * (size_t)p.cols * 32, tile_bytes,
                 bar_tile + (buf ^ 1) * 8);
      }
      const bool has_missing = __ldg(p.tile_flags + tile) != 0;
      mbar_wait(bar_tile + buf * 8, (tile_phase >> buf) & 1u);
      tile_phase ^= 1u << buf;

      const size_t row = tile * 32 + lane;
      const bool live = row < p.rows;
      float acc = (first || !live) ? 0.0f : p.preds[row];
      const uint8_t* tile_lane = smem + (my_tiles - s_base) + buf * tile_bytes + lane * 4;
      const uint32_t* forest = (const uint32_t*)(smem + (s_forest - s_base));

      int t = 0;
      for (; t + kGroup <= ntrees; t += kGroup) {
        uint32_t n[kGroup];
        const uint32_t* tree = forest + (uint32_t)t * W;
        if (has_missing) traverse_group<DEPTH, kGroup, true>(tree, tile_lane, p.col_mask, p.miss_code, n);
        else traverse_group<DEPTH, kGroup, false>(tree, tile_lane, p.col_mask, p.miss_code, n);
#pragma unroll
        for (int g = 0; g < kGroup; ++g) {
          acc += __uint_as_float(tree[g * W + n[g]]);
          if (LEAF_IDS && live)
            p.leaf_ids[row * (size_t)p.num_trees + tree0 + t + g] =
                __ldg(p.leaf_origin + (size_t)(tree0 + t + g) * L + (n[g] - L));
        }
      }
      for (; t < ntrees; ++t) {
        uint32_t n[1];
        const uint32_t* tree = forest + (uint32_t)t * W;
        if (has_missing) traverse_group<DEPTH, 1, true>(tree, tile_lane, p.col_mask, p.miss_code, n);
        else traverse_group<DEPTH, 1, false>(tree, tile_lane, p.col_mask, p.miss_code, n);
        acc += __uint_as_float(tree[n[0]]);
        if (LEAF_IDS && live)
          p.leaf_ids[row * (size_t)p.num_trees + tree0 + t] =
              __ldg(p.leaf_origin + (size_t)(tree0 + t) * L + (n[0] - L));
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
  if (p.leaf_ids) {
    TAHOE_CUDA(cudaFuncSetAttribute(k_smem_forest<DEPTH, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)f->smem_bytes));
    k_smem_forest<DEPTH, true><<<grid, kSmemThreads, f->smem_bytes, stream>>>(p);
  } else {
    TAHOE_CUDA(cudaFuncSetAttribute(k_smem_forest<DEPTH, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)f->smem_bytes));
    k_smem_forest<DEPTH, false><<<grid, kSmemThreads, f->smem_bytes, stream>>>(p);
  }
  TAHOE_CUDA(cudaGetLastError());
  return TAHOE_OK;
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
  q.table = f->d_qtable; q.offsets = f->d_qoffsets; q.rank_shift = f->rank_shift;
  q.miss_code = f->miss_code; q.qtiles = a.ws->qtiles; q.tile_flags = a.ws->tile_flags;
  q.num_tiles = tiles;
  const size_t qwarps_per_block = 256 / 32;
  unsigned qgrid = (unsigned)std::min<size_t>((tiles + qwarps_per_block - 1) / qwarps_per_block,
                                              (size_t)f->sm_count * 16);
  cudaEvent_t* ev = nullptr;
  if (f->profile && a.ws == &f->ws[0] && f->prof_count < kProfRing) {
    ev = &f->prof_events[(size_t)f->prof_count * 3];
    ++f->prof_count;
    TAHOE_CUDA(cudaEventRecord(ev[0], a.stream));
  }
  k_quantize<<<qgrid, 256, (q.cols + 1) * sizeof(int32_t), a.stream>>>(q);
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
