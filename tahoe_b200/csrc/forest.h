// forest.h -- internal representation behind the opaque tahoe_forest_t handle.
//
// One forest is packed into up to four layouts, each built for the memory level
// a strategy serves it from (DESIGN.md "Data layout in HBM"):
//
//   H8  heap layout, 8 B/node, tree-major, 1-based heap        DIRECT, PER_THREAD (L2 / HBM)
//   S16 compact layout, 16 B/node, breadth-first per tree      sparse forests      (L2 / HBM)
//   I8  node-major interleaved [node][tree], 8 B/node, children swapped by split
//       probability (exchange bit) and trees grouped by simhash  SHARED_DATA       (L1 / L2)
//   Q4  rank-quantised heap, 4 B/node, padded to full depth     SHARED_FOREST / SPLIT (shared memory)
//
// Reference counterparts: dense_adaptive_forest members bits_values_*/bits_*_reorg_*
// (Struct.h:2186-2207) and the packer that fills them (Struct.h:1756-1986).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/tahoe_b200.h"

namespace tahoe {

// k_smem_forest: levels 0 and 1 of every tree of a launch -- {root, left child, right child} -- travel as a
// kernel parameter (kernels_smem_traverse.cuh)
constexpr int kTopTrees = 2688;                 // 2688 * 12 B = 32,256 B (+ SmemParams < the 32,764-byte limit)
struct TopWords { uint32_t w[3 * kTopTrees]; };

constexpr int kMetaFidMask = (1 << 29) - 1;      // I8: fid in bits 0..28
constexpr uint32_t kMetaExchange = 1u << 29;     // I8 only (reference EXCHANGE_MASK_INT is bit 31, Struct.h:64)
constexpr uint32_t kMetaDefLeft = 1u << 30;      // same bit as the reference's DEF_LEFT_MASK, Struct.h:58
constexpr uint32_t kMetaLeaf = 1u << 31;         // same bit as IS_LEAF_MASK, Struct.h:59
// B32 (SHARED_DATA blocks, plan_shared_data): the node's second word as the walk wants it
constexpr uint32_t kSdMissRight = 1u << 31;      // a missing feature goes to the right CHILD SLOT (default direction and exchange folded)
constexpr uint32_t kSdFlip = 1u << 30;           // children exchanged: v >= thr goes to the LEFT child slot
constexpr uint32_t kSdLeaf = 1u << 29;           // pushed-down copy of a leaf (both children are the same leaf; feature offset 0)
constexpr uint32_t kSdOffMask = (1u << 29) - 1;  // byte offset of the feature inside the row

// Shared-memory-forest kernel: the plan (warps per CTA, tile buffering, trees per traversal group)
// is chosen per forest at pack time (plan_chunks) and is all run-time state of the handle; the
// environment variable TAHOE_SMEM_PLAN="warps,staged,group" overrides it for tuning runs.
constexpr int kSmemMaxWarps = 16;                // 512 threads, one persistent CTA per SM
constexpr int kMaxGroup = 6;                     // trees traversed together by one thread (ILP): 1..6
constexpr int kStagedMaxCols = 32;               // register-staged tiles keep one word per column per lane
constexpr int kStagedMaxColsDeep = 64;           // ... the sparse (deep) variant at <= 8 warps: 255 registers per thread
constexpr int kDeepPend = 16;                    // ring of parked deep walks per LANE (4 bytes each, power of two): k_smem_forest_deep
constexpr int ilog2_c(int v) { return v <= 1 ? 0 : 1 + ilog2_c(v >> 1); }

constexpr int kSparseWarps = 16;                 // k_sparse_smem: warps per persistent CTA
constexpr int kSparsePend = 160;                 // pending deep walks a warp can hold before it must drain
constexpr int kQuantColShift = 7;                // Q4 word, compact form: column index in bits 7.. (the wide
                                                 // form puts it at log2(128 * tile slots): the field is then
                                                 // the byte offset of the column row in the tile region)
constexpr int kQuantMaxDepth = 14;               // 2^(d+1) words per tree must fit shared memory
constexpr int kQuantMaxCols = 64;                // one warp tile = cols * 128 B of shared memory
constexpr int kQuantMaxBuckets = 8192;           // bucket index of one column: (B+1) ints of shared memory

struct HostPack {
  // H8 / S16
  std::vector<uint2> heap;            // dense: [T][2^(d+1)]; entry 0 of each tree unused
  std::vector<uint4> sparse;          // sparse: {val, bits, left_idx, 0}
  std::vector<uint2> sparse8;         // sparse, when the fields fit: {val, leaf<<31 | def_left<<30 | (left_idx - own index)<<fid_bits | fid}
  int sparse_fid_bits = 0;
  std::vector<int64_t> tree_offsets;  // sparse: [T+1]
  // T8 (sparse forests through shared memory): every tree renumbered breadth-first, 8 B/node as S8;
  // the first sp_cut nodes of each tree ("top") are what a chunk keeps in shared memory
  std::vector<uint2> sp_full;         // [total] breadth-first S8 nodes, same tree_offsets
  std::vector<int32_t> sp_orig;       // [total] breadth-first index -> caller's node index inside the tree
  std::vector<uint2> sp_top;          // [sp_chunks * sp_tpc][sp_cut]; short trees padded with leaf 0.0
  // QT (sparse forests, quantised tops): qwords / leaf_origin hold the dense top of depth qt_depth;
  // ... below them two-level blocks of 32 bytes {root, left, right, -} x {rank word | leaf value, leaf<<31 | block delta}
  std::vector<uint4> sp_qblocks;      // [blocks + 1][2]; the last one is the dummy leaf block (+0.0f)
  std::vector<int32_t> sp_qblock_orig;   // [blocks + 1][3] caller's node index of root / left / right (-1: none)
  std::vector<int64_t> sp_qblock_offsets;   // [T+1] first block of every tree
  // I8
  int tpad = 0;                       // trees padded to a multiple of 32
  std::vector<uint2> inter;           // [N][tpad]
  std::vector<int32_t> tree_order;    // [tpad] slot -> original tree, -1 padding
  std::vector<int32_t> orig_ids;      // [N][tpad] position -> original heap id
  std::vector<uint64_t> simhash;      // [T]
  // B32 (SHARED_DATA): leaves pushed down to the bottom level; the top sd_top levels as [node][tpad]
  // (shared-memory copy), the levels below in two-level blocks {root, left, right, -} of 32 bytes --
  // one L2 sector serves two levels of a walk
  std::vector<uint2> sd_top_nodes;    // [2^sd_top - 1][tpad]
  std::vector<uint2> sd_top_image;    // [2^sd_top - 1][sd_top_stride], padded to 16 bytes: what the kernel copies to shared memory
  std::vector<uint4> sd_blocks;       // [block][tpad][2]; block order: level sd_top roots, level sd_top+2 roots, ...
  std::vector<int32_t> sd_orig_bottom;  // [2^d][tpad] bottom position -> original heap id of the leaf reached
  // Q4
  std::vector<uint32_t> qwords;       // [T][2^(d+1)]
  std::vector<float> qtable;          // sorted distinct thresholds, all columns back to back
  std::vector<int32_t> qoffsets;      // [C+1]
  std::vector<int32_t> leaf_origin;   // [T][2^d]
  std::vector<uint32_t> top3;         // [qwords_trees][3] levels 0 and 1 of every (padded) tree: kernel-parameter table
  // bucket index over each column's threshold range: rank(v) is searched only inside v's bucket
  int qbuckets = 0;                   // B (power of two)
  std::vector<float> qlut_min;        // [C] smallest threshold of the column
  std::vector<float> qlut_scale;      // [C] B / (max - min)  (0 when the range is empty or infinite)
  std::vector<int32_t> qlut;          // [C][B+1] thresholds of the column whose bucket is < b
};

struct Workspace {                     // per-stream scratch, grow-only (no per-call malloc)
  uint32_t* qtiles = nullptr;          // [tiles][C][32] rank words
  uint32_t* tile_flags = nullptr;      // [tiles] bit0 = some feature of some row equals the sentinel
  size_t cap_tiles = 0;
  float* slab_data = nullptr;          // tahoe_predict_host staging
  float* slab_preds = nullptr;
  size_t cap_slab_rows = 0;
  float* partial = nullptr;            // tahoe_predict_to without caller scratch: partial sums between chunk passes
  size_t cap_partial = 0;
  cudaEvent_t done = nullptr;          // ws[0]: end of the last predict that used this scratch ...
  cudaStream_t last_stream = nullptr;  // ... and the stream it ran on (a call on another stream waits for it)
  bool used = false;
  cudaStream_t stream = nullptr;       // owned (predict_host only)
};

// Tuning knobs (measurement runs only): environment variables read ONCE, when the handle is created
// (query_device) -- nothing on the predict path calls getenv.
struct Tuning {
  size_t row_batch_bytes = 0;          // TAHOE_ROW_BATCH_BYTES: k_smem_forest walks the rows in L2-sized batches
  int sparse_grid = 0;                 // TAHOE_SPARSE_GRID: cap on k_sparse_smem's grid
  int sparse_group = 3;                // TAHOE_SPARSE_GROUP: sequences per lane of k_sparse_smem
  int direct_group = 0;                // TAHOE_DIRECT_GROUP: 4 = four walks per thread, 1 = one; 0 = by batch size
  bool no_reorder = false, no_simhash = false;   // TAHOE_NO_REORDER / TAHOE_NO_SIMHASH: ablation of the I8 layout (tools/probe_reorder.py)
  int cluster = -1;                    // TAHOE_CLUSTER: CTAs per cluster of the k_smem_forest cluster pipeline (0/1 = off, -1 = by plan)
  int quant_tiles = -1;                // TAHOE_QUANT_TILES: 1 = transpose to float tiles first, quantise them in place; 0 = from the rows; -1 = by table
  int top_levels = -1;                 // TAHOE_TOP_LEVELS: 0/1/2 tree levels from the kernel-parameter table; -1 = by plan
  bool sd_no_l2_ahead = false;         // TAHOE_SD_NO_L2_AHEAD: k_shared_data's single-buffer teams do not prefetch their next row into L2 (ablation)
  int top_trees = 0;                   // TAHOE_TOP_TREES: cap on the trees per launch of k_smem_forest (test knob: forces the split)
  int slab_first = 0, slab_growth = 0, slab_cap = 0;   // TAHOE_HOST_SLAB="first,growth%,cap" (waves of tiles); 0 = from the model
};

constexpr int kClassPad = 12;               // lcm of the traversal group sizes the sparse plans use (2, 4, 6) and of 1, 2, 3, 4, 6
constexpr int kHostSlots = 3;
constexpr int kProfRing = 512;         // per-kernel CUDA-event triples kept while profiling is on

}  // namespace tahoe

struct tahoe_forest {
  tahoe_forest_params_t params{};
  bool sparse = false;
  int depth = 0;                 // dense: given; sparse: max level found
  int64_t num_nodes = 0;         // real (reachable) nodes
  int node_width = 0;            // reference width rule
  int64_t exchanged = 0;
  double mean_depth = 0.0;

  // quantised layout parameters
  bool quant = false;
  int rank_bits = 0, col_bits = 0, rank_shift = 0, col_shift = 0;
  uint32_t col_mask = 0, miss_code = 0;
  int trees_per_chunk = 0, num_chunks = 0;
  int group = 1;                 // trees per traversal group (1..kMaxGroup); qwords_trees is a multiple
  int warps = tahoe::kSmemMaxWarps;  // warps per CTA of k_smem_forest
  bool staged = false;           // one tile buffer per warp, next tile staged in registers (else 2, by TMA)
  bool single = false;           // one tile buffer per warp filled by TMA after the walk, next tile prefetched into L2 (SINGLE)
  int slots = 0;                 // tile buffers per CTA = warps * (staged ? 1 : 2), a power of two
  // tile region of k_smem_forest: word (col, slot, lane) at col*col_stride + slot*slot_stride + lane*4
  // (wide: column-major, the column field of a node sits at bit 7 + log2 slots; compact: slot-major
  // with the columns rounded up to a power of two, column field at bit 7)
  uint32_t tile_col_stride = 0, tile_slot_stride = 0, tile_region = 0;
  int qwords_trees = 0;          // trees in d_qwords: num_trees padded with all-zero trees (leaf value +0.0f)
  size_t smem_bytes = 0;
  int quant_stride = 1;          // k_quantize keeps every quant_stride-th threshold of a column in shared memory
  size_t quant_smem = 0;

  // sparse shared-memory plan (plan_sparse_smem, at create): k_sparse_smem keeps the first sp_cut
  // breadth-first nodes of sp_tpc trees in shared memory; deeper nodes are walked in L2 (sp_full)
  int sp_cut = 0, sp_tpc = 0, sp_chunks = 0;
  size_t sp_smem = 0;
  double sp_overflow = 0.0;      // expected walks per (row, tree) that leave the top, children equally likely
  // QT plan (plan_sparse_qtop): the top qt_depth levels of every sparse tree as a dense Q4 heap served by
  // k_smem_forest<DEEP>; walks still on an internal node at that level are parked and finished in L2
  bool qt = false;
  bool sp_use_qt = false;        // QT is the path SHARED_FOREST / SPLIT take (else T8: TAHOE_SPARSE_PATH=t8 or a forced cut)
  int qt_depth = 0;
  double qt_overflow = 0.0;      // expected parked walks per (row, tree)
  int qdepth = 0;                // depth of the trees in d_qwords (dense: depth; sparse QT: qt_depth)

  // SHARED_DATA plan (plan_shared_data, at create): teams of team_warps warps share a row
  int sd_teams = 0, sd_team_warps = 1, sd_group = 1;
  int sd_row_bufs = 2;           // row buffers per team (1: single-buffered, twice the teams)
  int sd_top = 0;                // forest levels copied to shared memory
  bool sd_blocked = false;       // B32 two-level blocks below the top levels
  int sd_top_stride = 0;         // slots per node in the shared-memory copy of the top levels (tpad, or the tree count)
  int sd_per_sm[2] = {0, 0};     // CTAs per SM of the forest's k_shared_data launch, asked once ([0] LSU rows, [1] TMA rows)
  size_t sd_smem = 0;

  // per-class outputs (tahoe_forest_create_sparse_classes): class k owns packed trees
  // [class_first[k], class_first[k+1]) -- its real trees followed by single-leaf (+0.0f) padding trees up to a
  // multiple of kClassPad, so that a class is whole traversal groups under every plan -- class_real[k] of them real
  std::vector<int32_t> class_first, class_real;

  tahoe::HostPack host;
  tahoe::TopWords top_words_all{};   // host.top3 when the whole forest fits one launch's table

  // device copies
  uint2* d_heap = nullptr;
  uint4* d_sparse = nullptr;
  uint2* d_sparse8 = nullptr;
  uint2* d_sp_full = nullptr;
  int32_t* d_sp_orig = nullptr;
  uint2* d_sp_top = nullptr;
  uint4* d_sp_qblocks = nullptr;
  int32_t* d_sp_qblock_orig = nullptr;
  int64_t* d_sp_qblock_offsets = nullptr;
  int64_t* d_tree_offsets = nullptr;
  uint2* d_inter = nullptr;
  int32_t* d_tree_order = nullptr;
  int32_t* d_orig_ids = nullptr;
  uint2* d_sd_top_image = nullptr;
  uint4* d_sd_blocks = nullptr;
  int32_t* d_sd_orig_bottom = nullptr;
  uint32_t* d_qwords = nullptr;
  float* d_qtable = nullptr;
  int32_t* d_qoffsets = nullptr;
  int32_t* d_leaf_origin = nullptr;
  float* d_qlut_min = nullptr;
  float* d_qlut_scale = nullptr;
  int32_t* d_qlut = nullptr;

  int device = 0, sm_count = 0, smem_optin = 0, l2_bytes = 0;
  tahoe::Tuning tune;
  tahoe::Workspace ws[1 + tahoe::kHostSlots];  // [0] = tahoe_predict, [1..] = predict_host slabs
  int last_launches = 0;

  // optional per-kernel timing (tahoe_set_profiling): events recorded on the launch stream
  bool profile = false;
  int prof_count = 0;
  std::vector<cudaEvent_t> prof_events;  // 3 per entry: before quantise, between, after traverse
};

namespace tahoe {

void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what);
#define TAHOE_CUDA(call)                                         \
  do {                                                           \
    cudaError_t e__ = (call);                                    \
    if (e__ != cudaSuccess) return ::tahoe::cuda_fail(e__, #call); \
  } while (0)

// packer.cu
int pack_dense(tahoe_forest* f, const tahoe_dense_node_t* nodes);
int pack_sparse(tahoe_forest* f, const tahoe_sparse_node_t* nodes, const float* weights,
                const int64_t* tree_offsets);
int upload(tahoe_forest* f);
void plan_chunks(tahoe_forest* f);
void plan_shared_data(tahoe_forest* f);   // teams / top levels / B32 blocks; needs host.inter and query_device
double smem_visit_cost(int warps, int group, bool staged);   // relative cost of a node visit under a plan
double smem_tile_cost(int cols, int mode, int warps, int group, bool wide);   // one tile pass, in node visits (mode: 0 two TMA buffers, 1 staged, 2 SINGLE)
double smem_depth_factor(int depth);   // bank conflicts of the deep levels, relative to depth 10

// launchers (kernels_*.cu); each returns a status and adds to *launches
struct PredictArgs {
  const float* data;
  size_t rows;
  float* preds;        // partial sums between chunk passes (local memory; may be the same buffer as `out`)
  float* out;          // final scores, written once and never read back: may be peer-mapped
  int32_t* leaf_ids;
  cudaStream_t stream;
  Workspace* ws;
  // tahoe_predict_classes: only trees [tree_lo, tree_hi) of the packed forest (0, 0 = all of them), the AVG
  // transform divides by avg_trees (0 = params.num_trees), and the rank tiles of an earlier launch on the same
  // stream and workspace are reused (the quantise pass runs once per batch, not once per class)
  int tree_lo = 0, tree_hi = 0;
  int avg_trees = 0;
  bool reuse_tiles = false;
};
int launch_direct(tahoe_forest* f, const PredictArgs& a, bool rows_in_smem, int* launches);
int launch_shared_data(tahoe_forest* f, const PredictArgs& a, int* launches);
int launch_smem_forest(tahoe_forest* f, const PredictArgs& a, int* launches);
int smem_cluster_size(const tahoe_forest* f);   // CTAs per cluster of the shared-memory-forest plan (0 = plain kernel)
int launch_sparse_smem(tahoe_forest* f, const PredictArgs& a, int* launches);   // kernels_sparse_smem.cu
int launch_float_tiles(tahoe_forest* f, const float* data, size_t rows, float* tiles_out, uint32_t* tile_flags, size_t tiles,
                       cudaStream_t stream);   // kernels_sparse_smem.cu: rows -> column-major float tiles + missing flags
int query_device(tahoe_forest* f);
int ensure_workspace(tahoe_forest* f, Workspace* ws, size_t rows);

}  // namespace tahoe
