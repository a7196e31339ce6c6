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
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>

#include "kernels_smem_traverse.cuh"   // SmemParams; the kernel itself is instantiated per depth

namespace tahoe {

// ---------------------------------------------------------------------------------------------
struct QuantParams {
  const float* data;
  size_t rows;
  int cols;
  float missing;
  const float* table;      // sorted distinct thresholds, per column
  const int32_t* offsets;  // [cols] padded starts, then [cols] counts
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
// FROM_TILES: the values come from (and the rank words go back to) the column-major float tiles k_float_tiles wrote into
// the same buffer: a warp's 32 values are one 128-byte line instead of 32 sectors of the row-major matrix.
template <bool FROM_TILES>
__global__ void __launch_bounds__(kQuantThreads, 1) k_quantize(QuantParams p) {
  extern __shared__ __align__(16) uint8_t qsm[];
  const int c = blockIdx.x % p.cols;
  const int split = blockIdx.x / p.cols;
  const int B = p.buckets;
  const int S = 1 << p.stride_log2;
  const int off = p.offsets[c];                      // 16-byte aligned, NaN-padded (packer upload)
  const int n = p.offsets[p.cols + c];
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
    if (is_missing(v, p.missing)) {             // BaseTahoeTest.h:451-452 (+ NaN sentinel: device_utils.cuh)
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
    if (S > 1) {                                 // count inside that block: independent 16-byte loads
      const float4* blk = reinterpret_cast<const float4*>(tab + r);
      int cnt = 0;
      for (int j = 0; j < (S >> 2); ++j) {
        const float4 q = __ldg(blk + j);
        cnt += (q.x <= v) + (q.y <= v) + (q.z <= v) + (q.w <= v);
      }
      r += cnt;
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
      if (FROM_TILES) v[u] = live[u] ? __uint_as_float(__ldcg(p.qtiles + (tile * p.cols + c) * 32 + lane)) : 0.0f;
      else v[u] = live[u] ? __ldg(p.data + row * (size_t)p.cols + c) : 0.0f;
    }
#pragma unroll
    for (int u = 0; u < kQuantUnroll; ++u) {
      const size_t tile = base + (size_t)u * kWarps;
      if (tile >= t1) break;                     // warp-uniform
      bool miss = false;
      const uint32_t code = live[u] ? rank_code(v[u], miss) : 0u;
      p.qtiles[(tile * p.cols + c) * 32 + lane] = code;
      if (!FROM_TILES && __any_sync(0xffffffffu, miss) && lane == 0) atomicOr(p.tile_flags + tile, 1u);   // (k_float_tiles set the flags)
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

// one translation unit per depth (kernels_smem_depth.cu)
#define TAHOE_DEPTH_DECL(D) \
  int launch_smem_depth_##D(tahoe_forest* f, const SmemParams& p, const TopWords& tw, unsigned grid, cudaStream_t stream);
TAHOE_DEPTH_DECL(1) TAHOE_DEPTH_DECL(2) TAHOE_DEPTH_DECL(3) TAHOE_DEPTH_DECL(4) TAHOE_DEPTH_DECL(5)
TAHOE_DEPTH_DECL(6) TAHOE_DEPTH_DECL(7) TAHOE_DEPTH_DECL(8) TAHOE_DEPTH_DECL(9) TAHOE_DEPTH_DECL(10)
TAHOE_DEPTH_DECL(11) TAHOE_DEPTH_DECL(12) TAHOE_DEPTH_DECL(13) TAHOE_DEPTH_DECL(14)
#undef TAHOE_DEPTH_DECL

// Cluster pipeline (kernels_smem_cluster.cuh): CTAs per cluster the handle's plan runs with (0 = plain kernel).  Needs TMA
// tiles in the slot-major region with two spare column rows (the hand-over ring), one traversal group per chunk, a dense
// forest of depth >= 10 and enough chunks to share.
int smem_cluster_size(const tahoe_forest* f) {
  const bool able = !f->sparse && f->quant && !f->staged && !f->single && f->tile_col_stride == 128u &&
                    f->tile_slot_stride >= ((uint32_t)f->params.num_cols + 2u) * 128u &&
                    f->trees_per_chunk == f->group && f->group >= 2 && f->qdepth >= 10;
  // OPT-IN (TAHOE_CLUSTER=2).  Alone on the GPU, pairs of CTAs run config 3 at equal speed (52.9 vs 53.2 ms) with half the DRAM
  // traffic and are bit-exact in every test -- but through tahoe_predict_host, where the next slab's small kernels
  // (k_float_tiles) run CONCURRENTLY with the pipeline, repeated 10M-row calls came back with a few whole tiles wrong at
  // the slab boundaries (profiles/r02_cluster_pipeline.md, last section; the plain kernel is clean in the same runs).
  // Until that is understood the plain kernel is the default.
  const int S = std::min(f->tune.cluster >= 0 ? f->tune.cluster : 0, 8);
  return (able && S > 1 && f->num_chunks >= 2 * S) ? S : 0;
}

int launch_smem_forest(tahoe_forest* f, const PredictArgs& a, int* launches) {
  if (!(f->quant || f->qt) || f->trees_per_chunk < 1) {
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
  // Rows -> column-major float tiles first (k_float_tiles), then the ranks in place: pays where the column's threshold
  // table is strided (every value also gathers a 16-byte block of thresholds from L2, so the 32 row sectors per
  // warp-load hurt most): config 3 2.57 -> 2.17 ms, config 4 0.73 -> 0.55 ms; with the whole table in shared memory
  // the extra pass costs more than it saves (config 2: 0.42 -> 0.51 ms).  TAHOE_QUANT_TILES=0/1 overrides.
  const bool from_tiles = f->tune.quant_tiles >= 0 ? f->tune.quant_tiles > 0 : f->quant_stride > 1;
  if (f->quant_smem > 48 * 1024) {
    if (from_tiles) TAHOE_CUDA(cudaFuncSetAttribute(k_quantize<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f->quant_smem));
    else TAHOE_CUDA(cudaFuncSetAttribute(k_quantize<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f->quant_smem));
  }
  cudaEvent_t* ev = nullptr;
  if (f->profile && a.ws == &f->ws[0] && f->prof_count < kProfRing) {
    ev = &f->prof_events[(size_t)f->prof_count * 3];
    ++f->prof_count;
    TAHOE_CUDA(cudaEventRecord(ev[0], a.stream));
  }
  if (!a.reuse_tiles) {   // (a later class of tahoe_predict_classes walks the tiles its first class quantised)
    if (from_tiles) {
      // transpose first (coalesced both ways, sets the tile flags), then quantise the tiles in place
      int rc2 = launch_float_tiles(f, a.data, a.rows, reinterpret_cast<float*>(a.ws->qtiles), a.ws->tile_flags, tiles, a.stream);
      if (rc2) return rc2;
      k_quantize<true><<<qgrid, kQuantThreads, f->quant_smem, a.stream>>>(q);
      *launches += 2;
    } else {
      TAHOE_CUDA(cudaMemsetAsync(a.ws->tile_flags, 0, tiles * sizeof(uint32_t), a.stream));
      k_quantize<false><<<qgrid, kQuantThreads, f->quant_smem, a.stream>>>(q);
      ++*launches;
    }
  }
  if (ev) TAHOE_CUDA(cudaEventRecord(ev[1], a.stream));
  TAHOE_CUDA(cudaGetLastError());

  SmemParams p{};
  p.qtiles = a.ws->qtiles; p.tile_flags = a.ws->tile_flags; p.num_tiles = tiles; p.rows = a.rows;
  p.cols = f->params.num_cols; p.qwords = f->d_qwords; p.num_trees = f->params.num_trees;
  p.padded_trees = f->qwords_trees;
  p.trees_per_chunk = f->trees_per_chunk; p.num_chunks = f->num_chunks;
  p.col_mask = f->col_mask; p.miss_code = f->miss_code;
  // parameter-table levels pay when enough warps cover the predicated constant loads (config 2, 16 warps:
  // 8.23 -> 7.82 ms; config 4: 9.00 -> 8.84) and lose with fewer (config 3, 8 warps: 11.57 -> 12.0 ms on 2M rows; 12 warps: 11.01 -> 11.19)
  p.top_levels = f->tune.top_levels >= 0 ? (f->tune.top_levels >= 1 ? 2 : 0) : (f->warps >= 16 ? 2 : 0);
  p.col_stride = f->tile_col_stride; p.slot_stride = f->tile_slot_stride; p.tile_region = f->tile_region;
  p.preds = a.preds; p.out = a.out; p.leaf_ids = a.leaf_ids;
  p.leaf_origin = f->d_leaf_origin;
  p.deep_blocks = f->d_sp_qblocks; p.block_offsets = f->d_sp_qblock_offsets; p.block_orig = f->d_sp_qblock_orig;
  p.tf = Transform{f->params.output, (float)(a.avg_trees ? a.avg_trees : f->params.num_trees), f->params.threshold,
                   f->params.global_bias};
  p.cluster = p.leaf_ids ? 0 : smem_cluster_size(f);
  // Row batches (tuning knob, off by default): walk the rows in batches whose rank tiles stay
  // L2-resident between chunk passes.  Measured on B200 (profiles/r01_row_batch.log): never a win --
  // the tile stream hides behind the traversal, and every batch pays a pipeline fill and drain.
  size_t batch_tiles = tiles;
  if (f->tune.row_batch_bytes) {
    const size_t budget = f->tune.row_batch_bytes;
    const size_t wave = (size_t)f->sm_count * f->warps;
    const size_t fit = budget / ((size_t)p.cols * 128);
    if (budget && f->num_chunks > 1 && fit < tiles) batch_tiles = std::max<size_t>(wave, fit / wave * wave);
  }
  // Levels 0 and 1 of the trees ride in the kernel parameters (TopWords).  A forest with more trees than the
  // table holds is cut into consecutive chunk ranges, one launch each; partial sums flow from launch to
  // launch through preds[] exactly as they do from chunk to chunk.  (The leaf-id instantiation reads every
  // level from shared memory and is never split.)
  const size_t W = (size_t)1 << (f->qdepth + 1);
  const int top_limit = std::max(f->trees_per_chunk, std::min(kTopTrees, f->tune.top_trees > 0 ? f->tune.top_trees : kTopTrees));
  // the trees this call walks: all of them, or one class (whole traversal groups: kClassPad)
  const bool ranged = a.tree_hi > a.tree_lo;
  if (ranged && (p.leaf_ids || (a.tree_lo % f->group) || ((a.tree_hi - a.tree_lo) % f->group))) {
    set_error("tree ranges must be whole traversal groups and carry no leaf ids"); return TAHOE_ERR_INVALID;
  }
  const int range_lo = ranged ? a.tree_lo : 0;
  const int range_trees = ranged ? a.tree_hi - a.tree_lo : f->qwords_trees;
  const int range_chunks = ranged ? (range_trees + f->trees_per_chunk - 1) / f->trees_per_chunk : f->num_chunks;
  const int chunks_per_launch = p.leaf_ids ? range_chunks : std::max(1, top_limit / f->trees_per_chunk);
  if (f->trees_per_chunk > kTopTrees && !p.leaf_ids) { set_error("chunk larger than the top-level table"); return TAHOE_ERR_INVALID; }
  for (size_t t0 = 0; t0 < tiles && rc == TAHOE_OK; t0 += batch_tiles) {
    SmemParams b = p;
    b.num_tiles = std::min(batch_tiles, tiles - t0);
    b.rows = std::min(a.rows - t0 * 32, b.num_tiles * 32);
    b.qtiles = p.qtiles + t0 * (size_t)p.cols * 32;
    b.tile_flags = p.tile_flags + t0;
    b.preds = p.preds + t0 * 32;
    b.out = p.out + t0 * 32;
    if (p.leaf_ids) b.leaf_ids = p.leaf_ids + t0 * 32 * (size_t)p.num_trees;
    const unsigned grid = (unsigned)std::min<size_t>((size_t)f->sm_count,
                                                     (b.num_tiles + f->warps - 1) / f->warps);
    for (int c0 = 0; c0 < range_chunks && rc == TAHOE_OK; c0 += chunks_per_launch) {
      SmemParams r = b;
      r.num_chunks = std::min(chunks_per_launch, range_chunks - c0);
      r.tree_base = range_lo + c0 * f->trees_per_chunk;
      r.padded_trees = std::min(r.num_chunks * f->trees_per_chunk, range_trees - c0 * f->trees_per_chunk);
      r.qwords = p.qwords + (size_t)r.tree_base * W;
      r.first_pass = c0 == 0;
      r.last_pass = c0 + r.num_chunks >= range_chunks;
      const TopWords* tw = &f->top_words_all;           // whole forest in one launch: prepared at upload
      TopWords local;
      if (!p.leaf_ids && (r.tree_base != 0 || r.padded_trees != f->qwords_trees || f->qwords_trees > kTopTrees)) {
        std::memcpy(local.w, f->host.top3.data() + (size_t)r.tree_base * 3, (size_t)r.padded_trees * 3 * sizeof(uint32_t));
        tw = &local;
      }
      switch (f->qdepth) {
#define TAHOE_DEPTH_CASE(D) case D: rc = launch_smem_depth_##D(f, r, *tw, grid, a.stream); break;
        TAHOE_DEPTH_CASE(1) TAHOE_DEPTH_CASE(2) TAHOE_DEPTH_CASE(3) TAHOE_DEPTH_CASE(4)
        TAHOE_DEPTH_CASE(5) TAHOE_DEPTH_CASE(6) TAHOE_DEPTH_CASE(7) TAHOE_DEPTH_CASE(8)
        TAHOE_DEPTH_CASE(9) TAHOE_DEPTH_CASE(10) TAHOE_DEPTH_CASE(11) TAHOE_DEPTH_CASE(12)
        TAHOE_DEPTH_CASE(13) TAHOE_DEPTH_CASE(14)
#undef TAHOE_DEPTH_CASE
        default: set_error("unsupported depth"); return TAHOE_ERR_NOT_APPLICABLE;
      }
      ++*launches;
    }
  }
  if (rc) return rc;
  if (ev) TAHOE_CUDA(cudaEventRecord(ev[2], a.stream));
  return TAHOE_OK;
}

}  // namespace tahoe

// ---------------------------------------------------------------------------------------------
// tahoe_predict_no_alloc -- SPLIT_SHARED_FOREST / SHARED_FOREST (k_quantize + k_smem_forest) for callers that
// may not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No
// counterpart in the reference.
//
// The scratch the two kernels share -- the batch's rank tiles [tiles][cols][32] and one flag word per tile,
// which tahoe_predict keeps in a grow-only cudaMalloc'd workspace -- lives here in a file-scope __device__
// array of kNoAllocScratchBytes (16 MiB: every GPU that loads this library holds it until exit).  In host
// code the array's NAME is only a stand-in, not a device address: the address is asked for once per device,
// with that device current, through cudaGetSymbolAddress, and that pointer is what the kernels get.
// Every call shares the array and finds the previous call's contents in it: the flag words (which must start
// at zero) are cleared on the stream before k_quantize, the rank tiles are fully overwritten.  Calls that
// overlap on one GPU (two streams, two handles, two host threads) RACE on it.
namespace tahoe {

constexpr size_t kNoAllocScratchBytes = (size_t)TAHOE_NO_ALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_no_alloc_scratch[kNoAllocScratchBytes];

constexpr int kNoAllocMaxDevices = 64;
static void* g_no_alloc_ptr[kNoAllocMaxDevices];     // device address of the array, per device ordinal

// layout: [tiles] flag words, rounded up to 256 bytes, then [tiles][cols][32] rank words
static size_t no_alloc_flag_bytes(size_t tiles) { return (tiles * sizeof(uint32_t) + 255) / 256 * 256; }
static size_t no_alloc_max_tiles(int cols) {
  size_t t = kNoAllocScratchBytes / ((size_t)cols * 128 + sizeof(uint32_t));
  while (t > 0 && no_alloc_flag_bytes(t) + t * (size_t)cols * 128 > kNoAllocScratchBytes) --t;
  return t;
}

}  // namespace tahoe

extern "C" int tahoe_predict_no_alloc(tahoe_forest_t* f, const float* d_data, size_t num_rows, float* d_preds,
                                      void* stream) {
  using namespace tahoe;
  if (!f || (num_rows && (!d_data || !d_preds))) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  if (f->device < 0) { set_error("no CUDA device: this handle was packed on a GPU-less host and cannot predict"); return TAHOE_ERR_CUDA; }
  if (!(f->quant || (f->qt && f->sp_use_qt)) || f->trees_per_chunk < 1) {
    set_error("tahoe_predict_no_alloc runs the shared-memory-forest strategy: the forest needs the rank-quantised layout "
              "(tahoe_strategy_applicable(f, TAHOE_STRATEGY_SPLIT_SHARED_FOREST, rows))");
    return TAHOE_ERR_NOT_APPLICABLE;
  }
  if (num_rows == 0) return TAHOE_OK;
  const int cols = f->params.num_cols;
  const size_t tiles = (num_rows + 31) / 32;
  const size_t max_tiles = no_alloc_max_tiles(cols);
  if (tiles > max_tiles) {
    set_error("tahoe_predict_no_alloc: " + std::to_string(num_rows) + " rows of " + std::to_string(cols) +
              " columns need more than the " + std::to_string(kNoAllocScratchBytes >> 20) +
              " MiB static scratch; the limit for this forest is " + std::to_string(max_tiles * 32) +
              " rows per call (split the batch, or use tahoe_predict)");
    return TAHOE_ERR_INVALID;
  }
  int dev = 0;
  TAHOE_CUDA(cudaGetDevice(&dev));
  if (dev != f->device) { set_error("tahoe_predict_no_alloc: the handle's device is not the current device"); return TAHOE_ERR_INVALID; }
  if (dev >= kNoAllocMaxDevices) { set_error("tahoe_predict_no_alloc: device ordinal out of range"); return TAHOE_ERR_INVALID; }
  if (!g_no_alloc_ptr[dev]) {   // set-up, once per device, that device current: the array's address ON this device
    void* ptr = nullptr;
    TAHOE_CUDA(cudaGetSymbolAddress(&ptr, g_no_alloc_scratch));
    g_no_alloc_ptr[dev] = ptr;
  }
  unsigned char* base = static_cast<unsigned char*>(g_no_alloc_ptr[dev]);
  // A workspace that is already large enough: ensure_workspace() returns on its first line (tiles <= cap_tiles
  // was checked above), so nothing on this path calls cudaMalloc / cudaFree.  launch_smem_forest clears the
  // flag words with cudaMemsetAsync on the stream before k_quantize.  Not ws[0]: no profiling events.
  Workspace w{};
  w.tile_flags = reinterpret_cast<uint32_t*>(base);
  w.qtiles = reinterpret_cast<uint32_t*>(base + no_alloc_flag_bytes(tiles));
  w.cap_tiles = tiles;
  PredictArgs a{d_data, num_rows, d_preds, d_preds, nullptr, (cudaStream_t)stream, &w};
  int launches = 0;
  const int rc = launch_smem_forest(f, a, &launches);
  f->last_launches = launches;
  return rc;
}
