// abi.cu -- the extern "C" boundary declared in include/tahoe_b200.h.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <new>

#include "forest.h"

namespace tahoe {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
int cuda_fail(cudaError_t e, const char* what) {
  g_error = std::string(what) + ": " + cudaGetErrorString(e);
  cudaGetLastError();  // clear the sticky-free error state
  return TAHOE_ERR_CUDA;
}

static void free_workspace(Workspace& w) {
  if (w.qtiles) cudaFree(w.qtiles);
  if (w.tile_flags) cudaFree(w.tile_flags);
  if (w.slab_data) cudaFree(w.slab_data);
  if (w.slab_preds) cudaFree(w.slab_preds);
  if (w.partial) cudaFree(w.partial);
  if (w.done) cudaEventDestroy(w.done);
  if (w.stream) cudaStreamDestroy(w.stream);
  w = Workspace{};
}

static bool whole_forest_fits(const tahoe_forest* f) {
  if (f->sparse && f->sp_use_qt) return f->num_chunks == 1 && f->qt_overflow == 0.0;
  if (f->sparse) return f->sp_tpc >= 1 && f->sp_chunks == 1 && f->sp_overflow == 0.0;
  return f->quant && f->num_chunks == 1;
}

static int applicable(const tahoe_forest* f, int strategy, size_t rows) {
  (void)rows;
  switch (strategy) {
    case TAHOE_STRATEGY_DIRECT:
      return TAHOE_OK;
    case TAHOE_STRATEGY_SHARED_DATA_PER_THREAD:
      return (size_t)f->params.num_cols * 256 * 4 <= (size_t)f->smem_optin ? TAHOE_OK : TAHOE_ERR_NOT_APPLICABLE;
    case TAHOE_STRATEGY_SHARED_DATA:
      return (!f->sparse && !f->host.inter.empty() && f->sd_teams >= 1) ? TAHOE_OK : TAHOE_ERR_NOT_APPLICABLE;
    case TAHOE_STRATEGY_SHARED_FOREST:   // whole forest resident (BaseTahoeTest.h:657-668)
      return whole_forest_fits(f) ? TAHOE_OK : TAHOE_ERR_NOT_APPLICABLE;
    case TAHOE_STRATEGY_SPLIT_SHARED_FOREST:  // at least one tree per chunk (:670-679)
      if (f->sparse && f->sp_use_qt) return f->trees_per_chunk >= 1 ? TAHOE_OK : TAHOE_ERR_NOT_APPLICABLE;
      if (f->sparse) return f->sp_tpc >= 1 ? TAHOE_OK : TAHOE_ERR_NOT_APPLICABLE;
      return (f->quant && f->trees_per_chunk >= 1) ? TAHOE_OK : TAHOE_ERR_NOT_APPLICABLE;
    default:
      return TAHOE_ERR_INVALID;
  }
}

// Performance model (replaces main.cu:22-82).  Same idea as the reference -- per strategy, a time
// built from the forest/batch shape and a few machine rates, argmin wins -- but every strategy has a
// LATENCY term (what one warp's dependent chain costs when the batch cannot fill the machine) and a
// THROUGHPUT term (node visits over the rate the full machine sustains), and the rates are B200
// measurements (tools/sweep.py, profiles/r01_sweep_*.json), not multiples of a cudaMemcpy figure.
namespace model {
constexpr double kLaunch = 5e-6;            // launch + prologue of one kernel
constexpr double kL2Visit = 106e-9;         // one dependent L2 node visit of a thread walking 4 trees at once
constexpr double kDirectRate = 9.0e11;      // visits/s, DIRECT, forest far below L2 capacity
constexpr double kPerThreadRate = 1.3e12;   // visits/s, rows staged in shared memory
constexpr double kL1Bytes = 200e3;          // rows of the resident threads that L1 can hold
constexpr double kRowGather = 1.7e12;       // B/s of 32-byte row sectors gathered through L2
constexpr double kForestHalf = 80e6;        // forest bytes at which the L2-gather rates halve
constexpr double kTeamLevel = 0.7e-6;       // SHARED_DATA: one tree level of a row's walk (latency-bound)
constexpr double kSharedDataRate = 7.0e11;  // visits/s (slot-levels), SHARED_DATA
constexpr double kTilePass = 4.6e12;        // B/s through the tile pass of the 8-byte sparse tops (read + write)
constexpr double kRowStream = 6.0e12;       // B/s of sample rows through a TMA-fed SHARED_DATA kernel (config 1: 614 MB in 0.109 ms at 50K rows, 2.46 GB in 0.364 ms at 200K)
// The shared-memory forest's rate is DERIVED FROM NCU COUNTERS of the unit plan (config 2, 16 warps x 5 trees, SINGLE
// tiles, parameter-table levels; profiles/r02_ncu_full_C2_summary.json), not fitted to wall-clock:
//   l1tex__data_pipe_lsu_wavefronts_mem_shared.sum = 2.0294e9 over rows x trees x depth = 5M x 500 x 10 lane-visits
//   l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed = 95.6 %; the pipe retires one wavefront per clock
//   per SM (tahoe_measure_peaks: 37.1 TB/s = 128 B x 148 SMs x 1.965 GHz)
// => seconds per lane-visit = wavefronts per lane-visit / (pipe rate x utilisation) = 2.92e-13.
// The same two counters on config 3 (1.0898e10 wavefronts over 1.2e11 lane-visits at 89.4 %) give 3.49e-13 before the
// parameter-table discount both share: the ratio, 1.20 in counters and 1.3 in wall-clock (46.5 ms), is the planner's
// smem_depth_factor(12) = 1.3.
constexpr double kSmemWavefrontsPerLaneVisit = 2.0294e9 / 2.5e10;
constexpr double kLsuPipeRate = 148.0 * 1.965e9;   // wavefronts/s
constexpr double kLsuPipeUtil = 0.956;
constexpr double kSmemVisit = kSmemWavefrontsPerLaneVisit / (kLsuPipeRate * kLsuPipeUtil);   // s per lane-visit, unit plan
constexpr double kChunkFixed = 4.0e-6;      // stage one chunk + barrier + the ragged end of a pass (config 3: 0.9 ms over 167 chunks)
constexpr double kTileLevel = 30e-9;        // one level of one traversal group of a tile: this much alone ...
constexpr double kTileLevelPerWarp = 5e-9;    // ... plus this per warp sharing the SM (16 warps: 110 ns, 8: 70 ns)
constexpr double kQuantRate = 1.7e12;       // B/s in+out of k_quantize, whole table in shared memory
constexpr double kQuantRateStrided = 1.0e12;   // ... every 2^k-th threshold there, rest probed in L2
constexpr double kSparseVisit = 2.3e-12;      // s per node visit, sparse tree tops in shared memory (free-running lanes)
constexpr double kSparseChunkRow = 1.0e-10;   // s per row per chunk: tile re-stream + refill latency share
constexpr double kSparseDeep = 3.1e-11;       // s per walk that leaves the tops and is finished in L2
constexpr double kSparseStep = 300e-9;        // one step of a lane's two interleaved sequences (latency)
constexpr double kDeepWalk = 1.3e-11;          // s per parked walk of the QT layout (two-level blocks in L2, pipelined behind the traversal)
constexpr double kSparseTileFixed = 8e-6;     // refill a warp's tile buffer + final drain
constexpr double kHostCopyRate = 50e9;        // B/s, pinned host -> device over PCIe Gen5 x16 (measured 57 GB/s, one GPU active)
}  // namespace model

static double forest_bytes(const tahoe_forest* f) {
  // what the full-machine DIRECT kernels gather from: the 8-byte compact nodes when the fields fit (k_sparse8;
  // config 4: 115 MB, inside L2 -- the 16-byte form is 229 MB and made the model predict 80.6 ms where 34.4 was measured)
  if (f->sparse) return f->host.sparse8.empty() ? (double)f->host.sparse.size() * 16.0 : (double)f->host.sparse8.size() * 8.0;
  return (double)f->host.heap.size() * 8.0;
}

static int select_strategy(const tahoe_forest* f, size_t rows_, double* out) {
  using namespace model;
  const double R = (double)std::max<size_t>(rows_, 1);
  const double T = f->params.num_trees, C = f->params.num_cols;
  const double dbar = std::max(f->mean_depth, 1.0);
  // a thread's walk through one tree ends with the warp's deepest lane: sparse trees are uneven
  const double dwalk = f->sparse ? 0.5 * (dbar + f->depth) : dbar;
  const double l2_scale = 1.0 / (1.0 + forest_bytes(f) / kForestHalf);
  int best = TAHOE_STRATEGY_DIRECT;
  double best_t = std::numeric_limits<double>::infinity();
  for (int s = 0; s < TAHOE_NUM_STRATEGIES; ++s) {
    double t = std::numeric_limits<double>::infinity();
    if (applicable(f, s, rows_) == TAHOE_OK) {
      switch (s) {
        case TAHOE_STRATEGY_DIRECT:
        case TAHOE_STRATEGY_SHARED_DATA_PER_THREAD: {
          const bool staged = s == TAHOE_STRATEGY_SHARED_DATA_PER_THREAD;
          // sparse trees are uneven: a warp's lanes idle until its deepest walk ends (14 of 32 lanes
          // active on config 4), and staging rows does not pay for itself there
          const double rate = (f->sparse ? kDirectRate / (staged ? 1.65 : 1.25) : (staged ? kPerThreadRate : kDirectRate)) * l2_scale;
          const double lat = T * dwalk * kL2Visit * (staged ? 0.96 : 1.0);
          double thr = R * T * dbar / rate;
          // DIRECT gathers features straight from global memory: once the resident threads' rows
          // outgrow L1, every visit pulls a 32-byte sector of its row through L2
          // (the share of visits that miss grows with how far the rows overflow L1)
          if (!staged && C * 4.0 * 2048.0 > kL1Bytes) {
            const double miss = std::min(1.0, (C * 4.0 * 2048.0 / kL1Bytes - 1.0) / 8.0);
            // (sparse walks are short and end at different levels: fewer of a warp's lanes gather at once -- 0.4 fits
            // config 4: 34.4 ms measured at 1M rows, 10.0 at 300K)
            thr += (f->sparse ? 0.4 : 1.0) * miss * R * T * dbar * 32.0 / kRowGather;
          }
          // one thread per row: below a full wave of threads the chain and the machine's rate are both felt, so the
          // two terms ADD (held-out forest 50 x d14 x 28 at 100K rows: 0.186 ms measured, max() said 0.079, the sum 0.141)
          t = kLaunch + lat + thr;
          break;
        }
        case TAHOE_STRATEGY_SHARED_DATA: {
          const double tpad = f->host.tpad;
          const double lat = (f->depth + 1) * kTeamLevel;
          const double thr = R * tpad * dbar / kSharedDataRate;
          t = kLaunch + std::max(std::max(lat, thr), R * C * 4.0 / kRowStream);
          break;
        }
        default: if (f->sparse && !f->sp_use_qt) {   // T8: tile pass + tree tops in shared memory + deep walks finished in L2
          const double warps = (double)f->sm_count * kSparseWarps;
          const double tiles_per_warp = std::ceil(R / 32.0 / warps);
          const double visits = R * T * (dbar + 1.0);
          const double thr = visits * kSparseVisit + f->sp_chunks * R * kSparseChunkRow +
                             R * T * f->sp_overflow * kSparseDeep;
          // one warp's chain: two sequences of tpc/2 trees per tile, then the tile refill
          const double lat = f->sp_chunks * (kChunkFixed + tiles_per_warp *
                             (0.5 * f->sp_tpc * (dbar + 1.0) * kSparseStep + kSparseTileFixed));
          t = 2.0 * kLaunch + R * C * 8.0 / kTilePass + std::max(lat, thr);
          break;
        } else {   // SHARED_FOREST / SPLIT_SHARED_FOREST: quantise pass + chunk passes
          // (sparse forests, QT layout: the same kernel on the dense tops + parked deep walks in L2)
          const int depth = f->qdepth;
          const double plan = smem_visit_cost(f->warps, f->group, f->staged) * (f->sparse ? 1.0 : smem_depth_factor(depth)) *
                              ((f->warps >= 16 && !f->sparse) ? 1.0 - 0.25 * std::min(2, depth) / std::max(depth, 1) : 1.0);
          const bool wide = f->tile_col_stride != 128u;
          const double visits = R * (double)f->qwords_trees * depth;               // padded depth
          const double tiles_per_warp = std::ceil(R / 32.0 / ((double)f->sm_count * f->warps));
          const double groups = (double)f->trees_per_chunk / std::max(f->group, 1);
          const double lat = f->num_chunks * (kChunkFixed + tiles_per_warp * groups * depth *
                                              (kTileLevel + kTileLevelPerWarp * f->warps));
          const double thr = kSmemVisit * (visits * plan + f->num_chunks * R *
                             smem_tile_cost(f->params.num_cols, f->single ? 2 : (f->staged ? 1 : 0), f->warps, f->group, wide)) +
                             (f->sparse ? R * T * f->qt_overflow * kDeepWalk : 0.0);
          const double quant = kLaunch + R * C * 8.0 / (f->quant_stride > 1 ? kQuantRateStrided : kQuantRate);
          t = kLaunch + quant + std::max(lat, thr);
          break;
        }
      }
    }
    if (out) out[s] = t;
    if (t < best_t) { best_t = t; best = s; }
  }
  return best;
}

static int predict_impl(tahoe_forest* f, int strategy, const PredictArgs& a) {
  int launches = 0;
  if (f->device < 0) {
    set_error("no CUDA device: this handle was packed on a GPU-less host and cannot predict");
    return TAHOE_ERR_CUDA;
  }
  if (strategy == TAHOE_STRATEGY_AUTO) strategy = select_strategy(f, a.rows, nullptr);
  int rc = applicable(f, strategy, a.rows);
  if (rc == TAHOE_ERR_INVALID) { set_error("unknown strategy id"); return rc; }
  if (rc != TAHOE_OK) { set_error("Strategy is not suitable for this case."); return rc; }
  // profiling: the single-kernel strategies are bracketed here (no preparation pass: ev[0] == ev[1]);
  // the shared-memory-forest launchers record around their two kernels themselves
  cudaEvent_t* ev = nullptr;
  const bool one_kernel = strategy == TAHOE_STRATEGY_DIRECT || strategy == TAHOE_STRATEGY_SHARED_DATA_PER_THREAD ||
                          strategy == TAHOE_STRATEGY_SHARED_DATA;
  if (one_kernel && a.rows && f->profile && a.ws == &f->ws[0] && f->prof_count < kProfRing) {
    ev = &f->prof_events[(size_t)f->prof_count * 3];
    ++f->prof_count;
    TAHOE_CUDA(cudaEventRecord(ev[0], a.stream));
    TAHOE_CUDA(cudaEventRecord(ev[1], a.stream));
  }
  switch (strategy) {
    case TAHOE_STRATEGY_DIRECT: rc = launch_direct(f, a, false, &launches); break;
    case TAHOE_STRATEGY_SHARED_DATA_PER_THREAD: rc = launch_direct(f, a, true, &launches); break;
    case TAHOE_STRATEGY_SHARED_DATA: rc = launch_shared_data(f, a, &launches); break;
    default: rc = (f->sparse && !f->sp_use_qt) ? launch_sparse_smem(f, a, &launches) : launch_smem_forest(f, a, &launches); break;
  }
  if (ev) TAHOE_CUDA(cudaEventRecord(ev[2], a.stream));
  f->last_launches = launches;
  return rc;
}

}  // namespace tahoe

using namespace tahoe;

extern "C" {

const char* tahoe_last_error(void) { return g_error.c_str(); }
const char* tahoe_version(void) { return "tahoe_b200 0.1 (sm_100a)"; }

int tahoe_forest_create_dense(const tahoe_dense_node_t* nodes, const tahoe_forest_params_t* params,
                              tahoe_forest_t** out) {
  if (!nodes || !params || !out) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  *out = nullptr;
  tahoe_forest* f = new (std::nothrow) tahoe_forest;
  if (!f) return TAHOE_ERR_NOMEM;
  f->params = *params;
  int rc = query_device(f);
  try {
    if (rc == TAHOE_OK) rc = pack_dense(f, nodes);
  } catch (const std::bad_alloc&) {
    set_error("out of host memory while packing"); rc = TAHOE_ERR_NOMEM;
  }
  if (rc == TAHOE_OK) rc = upload(f);
  if (rc != TAHOE_OK) { tahoe_forest_destroy(f); return rc; }
  *out = f;
  return TAHOE_OK;
}

int tahoe_forest_create_sparse(const tahoe_sparse_node_t* nodes, const float* weights,
                               const int64_t* tree_offsets, const tahoe_forest_params_t* params,
                               tahoe_forest_t** out) {
  if (!nodes || !params || !out || !tree_offsets) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  *out = nullptr;
  tahoe_forest* f = new (std::nothrow) tahoe_forest;
  if (!f) return TAHOE_ERR_NOMEM;
  f->params = *params;
  int rc = query_device(f);
  try {
    if (rc == TAHOE_OK) rc = pack_sparse(f, nodes, weights, tree_offsets);
  } catch (const std::bad_alloc&) {
    set_error("out of host memory while packing"); rc = TAHOE_ERR_NOMEM;
  }
  if (rc == TAHOE_OK) rc = upload(f);
  if (rc != TAHOE_OK) { tahoe_forest_destroy(f); return rc; }
  *out = f;
  return TAHOE_OK;
}

static int predict_on_ws0(tahoe_forest* f, int strategy, PredictArgs a);   // below, with tahoe_predict

int tahoe_forest_create_sparse_classes(const tahoe_sparse_node_t* nodes, const float* weights,
                                       const int64_t* tree_offsets, const tahoe_forest_params_t* params,
                                       int num_classes, const int32_t* tree_class, tahoe_forest_t** out) {
  if (!nodes || !params || !out || !tree_offsets || !tree_class) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  *out = nullptr;
  const int T = params->num_trees;
  if (num_classes < 1 || T < 1) { set_error("need at least one class and one tree"); return TAHOE_ERR_INVALID; }
  std::vector<int32_t> class_first((size_t)num_classes + 1, 0), class_real((size_t)num_classes, 0);
  for (int t = 0; t < T; ++t) {
    if (tree_class[t] < 0 || tree_class[t] >= num_classes) { set_error("tree_class out of range"); return TAHOE_ERR_INVALID; }
    ++class_real[(size_t)tree_class[t]];
  }
  // trees grouped by class (stable: training order inside a class), every class padded to a multiple of kClassPad
  // with single-leaf trees worth +0.0f, so that a class is whole traversal groups under every plan
  try {
    std::vector<tahoe_sparse_node_t> pn;
    std::vector<float> pw;
    std::vector<int64_t> po(1, 0);
    for (int k = 0; k < num_classes; ++k) {
      class_first[(size_t)k] = (int32_t)po.size() - 1;
      for (int t = 0; t < T; ++t) {
        if (tree_class[t] != k) continue;
        const int64_t b = tree_offsets[t], e = tree_offsets[t + 1];
        if (e <= b) { set_error("empty tree"); return TAHOE_ERR_INVALID; }
        pn.insert(pn.end(), nodes + b, nodes + e);
        if (weights) pw.insert(pw.end(), weights + b, weights + e);
        po.push_back((int64_t)pn.size());
      }
      while ((((int64_t)po.size() - 1) - class_first[(size_t)k]) % kClassPad) {
        pn.push_back(tahoe_sparse_node_t{0.0f, (int32_t)0x80000000, 0});
        if (weights) pw.push_back(1.0f);
        po.push_back((int64_t)pn.size());
      }
    }
    class_first[(size_t)num_classes] = (int32_t)po.size() - 1;
    tahoe_forest_params_t pp = *params;
    pp.num_trees = (int)po.size() - 1;
    pp.num_nodes = (int)pn.size();
    const int rc = tahoe_forest_create_sparse(pn.data(), weights ? pw.data() : nullptr, po.data(), &pp, out);
    if (rc != TAHOE_OK) return rc;
    (*out)->class_first = class_first;
    (*out)->class_real = class_real;
  } catch (const std::bad_alloc&) {
    set_error("out of host memory while packing"); return TAHOE_ERR_NOMEM;
  }
  return TAHOE_OK;
}

int tahoe_predict_classes(tahoe_forest_t* f, int strategy, const float* d_data, size_t num_rows, float* d_out,
                          void* stream) {
  if (!f || (num_rows && (!d_data || !d_out))) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  if (f->class_first.empty()) { set_error("not a class forest: create it with tahoe_forest_create_sparse_classes"); return TAHOE_ERR_INVALID; }
  if (strategy == TAHOE_STRATEGY_AUTO) strategy = select_strategy(f, num_rows, nullptr);
  const bool smem = strategy == TAHOE_STRATEGY_SHARED_FOREST || strategy == TAHOE_STRATEGY_SPLIT_SHARED_FOREST;
  // class ranges need the layouts that address trees one by one: the compact nodes (DIRECT / PER_THREAD) and the
  // quantised tops (shared-memory forest); anything else falls back to DIRECT
  if ((smem && !(f->qt && f->sp_use_qt)) || strategy == TAHOE_STRATEGY_SHARED_DATA) strategy = TAHOE_STRATEGY_DIRECT;
  const int K = (int)f->class_real.size();
  int total = 0;
  for (int k = 0; k < K; ++k) {
    PredictArgs a{d_data, num_rows, d_out + (size_t)k * num_rows, d_out + (size_t)k * num_rows, nullptr,
                  (cudaStream_t)stream, &f->ws[0]};
    a.tree_lo = f->class_first[(size_t)k];
    a.tree_hi = f->class_first[(size_t)k + 1];
    a.avg_trees = std::max(1, f->class_real[(size_t)k]);
    a.reuse_tiles = smem && k > 0;        // ONE quantise pass for the K classes
    // every class is a whole forest of its own to the kernels: first and last chunk pass inside its range
    const int rc = k == 0 ? predict_on_ws0(f, strategy, a) : predict_impl(f, strategy, a);
    if (rc != TAHOE_OK) return rc;
    total += f->last_launches;
  }
  if (f->device >= 0 && smem) {   // the event predict_on_ws0 recorded covered class 0 only (the other strategies keep
    Workspace& w = f->ws[0];      // nothing in the workspace: no event)
    TAHOE_CUDA(cudaEventRecord(w.done, (cudaStream_t)stream));
  }
  f->last_launches = total;
  return TAHOE_OK;
}

void tahoe_forest_destroy(tahoe_forest_t* f) {
  if (!f) return;
  cudaFree(f->d_heap); cudaFree(f->d_sparse); cudaFree(f->d_sparse8); cudaFree(f->d_tree_offsets); cudaFree(f->d_inter);
  cudaFree(f->d_tree_order); cudaFree(f->d_orig_ids);
  cudaFree(f->d_sp_full); cudaFree(f->d_sp_orig); cudaFree(f->d_sp_top); cudaFree(f->d_sp_qblocks);
  cudaFree(f->d_sp_qblock_orig); cudaFree(f->d_sp_qblock_offsets);
  cudaFree(f->d_sd_top_image); cudaFree(f->d_sd_blocks); cudaFree(f->d_sd_orig_bottom); cudaFree(f->d_qwords); cudaFree(f->d_qtable);
  cudaFree(f->d_qoffsets); cudaFree(f->d_leaf_origin);
  cudaFree(f->d_qlut_min); cudaFree(f->d_qlut_scale); cudaFree(f->d_qlut);
  for (auto& w : f->ws) free_workspace(w);
  for (auto& e : f->prof_events) cudaEventDestroy(e);
  cudaGetLastError();
  delete f;
}

int tahoe_forest_info(const tahoe_forest_t* f, tahoe_forest_info_t* info) {
  if (!f || !info) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  std::memset(info, 0, sizeof(*info));
  info->num_trees = f->params.num_trees; info->depth = f->depth; info->num_cols = f->params.num_cols;
  info->is_sparse = f->sparse; info->num_nodes = f->num_nodes; info->node_width_bytes = f->node_width;
  info->quantized = f->quant; info->rank_bits = f->rank_bits; info->col_bits = f->col_bits;
  info->trees_per_chunk = f->sparse ? f->sp_tpc : f->trees_per_chunk;
  info->num_chunks = f->sparse ? f->sp_chunks : f->num_chunks;
  info->sparse_cut = f->sp_cut; info->sparse_deep_walks = f->sp_overflow;
  info->sparse_fid_bits = f->host.sparse_fid_bits;
  info->sparse_top_depth = f->qt ? f->qt_depth : 0; info->sparse_top_active = f->sp_use_qt;
  info->sparse_top_trees_per_chunk = f->qt ? f->trees_per_chunk : 0; info->sparse_top_chunks = f->qt ? f->num_chunks : 0;
  info->sparse_top_deep_walks = f->qt_overflow;
  info->smem_cluster = smem_cluster_size(f);
  info->bytes_heap = (int64_t)(f->host.heap.size() * sizeof(uint2) + f->host.sparse.size() * sizeof(uint4));
  info->bytes_interleaved = (int64_t)(f->host.inter.size() * sizeof(uint2));
  info->bytes_quantized = (int64_t)(f->host.qwords.size() * sizeof(uint32_t));
  info->exchanged_nodes = f->exchanged; info->mean_depth = f->mean_depth; info->device = f->device;
  info->sm_count = f->sm_count; info->smem_optin_bytes = f->smem_optin; info->col_shift = f->col_shift;
  info->smem_warps = f->warps; info->smem_staged = f->single ? 2 : (f->staged ? 1 : 0); info->smem_group = f->group;
  info->sd_teams = f->sd_teams; info->sd_team_warps = f->sd_team_warps;
  info->sd_top_levels = f->sd_top; info->sd_blocked = f->sd_blocked;
  return TAHOE_OK;
}

// tahoe_predict / tahoe_predict_to share the handle's workspace ws[0] (rank tiles, tile flags, partial sums).
// Two calls on DIFFERENT streams must not overlap on it: the second stream waits for the event the first
// call recorded (stream-ordered, no host sync).  Calls on one stream are ordered by the stream itself.
static int predict_on_ws0(tahoe_forest* f, int strategy, PredictArgs a) {
  Workspace& w = f->ws[0];
  if (strategy == TAHOE_STRATEGY_AUTO && f->device >= 0) strategy = select_strategy(f, a.rows, nullptr);
  // the one-kernel strategies keep nothing in the handle's workspace: no event traffic on their launch path
  // (small batches are bound by the host side of the call)
  const bool uses_ws = !(strategy == TAHOE_STRATEGY_DIRECT || strategy == TAHOE_STRATEGY_SHARED_DATA_PER_THREAD ||
                         strategy == TAHOE_STRATEGY_SHARED_DATA);
  if (f->device >= 0 && uses_ws) {
    if (!w.done) TAHOE_CUDA(cudaEventCreateWithFlags(&w.done, cudaEventDisableTiming));
    if (w.used && w.last_stream != a.stream) TAHOE_CUDA(cudaStreamWaitEvent(a.stream, w.done, 0));
  }
  const int rc = predict_impl(f, strategy, a);
  if (f->device >= 0 && rc == TAHOE_OK && uses_ws) {
    TAHOE_CUDA(cudaEventRecord(w.done, a.stream));
    w.used = true; w.last_stream = a.stream;
  }
  return rc;
}

int tahoe_predict(tahoe_forest_t* f, int strategy, const float* d_data, size_t num_rows,
                  float* d_preds, int32_t* d_leaf_ids, void* stream) {
  if (!f || (num_rows && (!d_data || !d_preds))) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  PredictArgs a{d_data, num_rows, d_preds, d_preds, d_leaf_ids, (cudaStream_t)stream, &f->ws[0]};
  return predict_on_ws0(f, strategy, a);
}

int tahoe_predict_to(tahoe_forest_t* f, int strategy, const float* d_data, size_t num_rows,
                     float* d_out, float* d_partial, int32_t* d_leaf_ids, void* stream) {
  if (!f || (num_rows && (!d_data || !d_out))) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  if (!d_partial && num_rows && f->device >= 0) {   // handle-owned scratch, grow-only
    Workspace& w = f->ws[0];
    if (w.cap_partial < num_rows) {
      // (a grow while an earlier call still runs on another stream: wait for it before freeing)
      if (w.used) TAHOE_CUDA(cudaEventSynchronize(w.done));
      if (w.partial) cudaFree(w.partial);
      w.partial = nullptr; w.cap_partial = 0;
      TAHOE_CUDA(cudaMalloc((void**)&w.partial, num_rows * sizeof(float)));
      w.cap_partial = num_rows;
    }
    d_partial = w.partial;
  }
  PredictArgs a{d_data, num_rows, d_partial, d_out, d_leaf_ids, (cudaStream_t)stream, &f->ws[0]};
  return predict_on_ws0(f, strategy, a);
}

int tahoe_gather_buffer_create(size_t num_floats, float** d_ptr, unsigned char* ipc_handle) {
  if (!d_ptr || !ipc_handle) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  static_assert(sizeof(cudaIpcMemHandle_t) == TAHOE_IPC_HANDLE_BYTES, "IPC handle size");
  *d_ptr = nullptr;
  // its own cudaMalloc (not a slice of a caching allocator's block): the IPC handle then names exactly this buffer
  TAHOE_CUDA(cudaMalloc((void**)d_ptr, std::max<size_t>(num_floats, 1) * sizeof(float)));
  cudaIpcMemHandle_t h;
  const cudaError_t e = cudaIpcGetMemHandle(&h, *d_ptr);
  if (e != cudaSuccess) { cudaFree(*d_ptr); *d_ptr = nullptr; return cuda_fail(e, "cudaIpcGetMemHandle"); }
  std::memcpy(ipc_handle, &h, sizeof(h));
  return TAHOE_OK;
}

int tahoe_gather_buffer_open(const unsigned char* ipc_handle, float** d_ptr) {
  if (!d_ptr || !ipc_handle) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  *d_ptr = nullptr;
  cudaIpcMemHandle_t h;
  std::memcpy(&h, ipc_handle, sizeof(h));
  // opened with the CONSUMER's device current: the lazy flag maps the owner's memory into this device's address
  // space over NVLink.  (A mapping made in the owner device's context -- what torch's CUDA IPC does -- cannot be
  // stored to by kernels that run on another GPU: illegal address, found on two B200s.)
  TAHOE_CUDA(cudaIpcOpenMemHandle((void**)d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return TAHOE_OK;
}

int tahoe_gather_buffer_release(float* d_ptr, int owner) {
  if (!d_ptr) return TAHOE_OK;
  TAHOE_CUDA(owner ? cudaFree(d_ptr) : cudaIpcCloseMemHandle(d_ptr));
  return TAHOE_OK;
}

int tahoe_enable_peer_access(int peer_device) {
  int cur = 0;
  TAHOE_CUDA(cudaGetDevice(&cur));
  if (cur == peer_device) return TAHOE_OK;
  int can = 0;
  TAHOE_CUDA(cudaDeviceCanAccessPeer(&can, cur, peer_device));
  if (!can) { set_error("the current device cannot access the peer device's memory"); return TAHOE_ERR_NOT_APPLICABLE; }
  const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return TAHOE_OK; }
  if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceEnablePeerAccess");
  return TAHOE_OK;
}

static int predict_host_impl(tahoe_forest_t* f, int strategy, const float* h_data, size_t num_rows,
                             float* h_preds, bool copy_only) {
  if (!f || (num_rows && (!h_data || !h_preds))) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  if (f->device < 0) { set_error("no CUDA device: this handle was packed on a GPU-less host and cannot predict"); return TAHOE_ERR_CUDA; }
  if (num_rows == 0) return TAHOE_OK;
  const int C = f->params.num_cols;
  if (strategy == TAHOE_STRATEGY_AUTO) strategy = select_strategy(f, num_rows, nullptr);
  // Slabs: copy of slab i+1 under the kernel of slab i, three slabs in flight.  The shared-memory
  // strategies deal 32-row tiles to sm_count x warps persistent warps, so their slabs are whole WAVES
  // of tiles (no ragged last round per chunk pass).  The first slab is short (2 waves: its copy is
  // the one nothing can hide) and every slab is `growth` times the one before, up to `cap` waves: a slab may
  // only grow as fast as the kernels before it run ahead of the copy engine, so growth follows the
  // ratio of copy time to predicted kernel time per row -- 1.5x / 24 waves where the copy takes 0.8 of the
  // predict (config 2), 3x / 96 waves where it takes 0.37 (config 3: every slab kernel pays 167 chunk
  // passes of fixed cost, so fewer, larger slabs win) -- tools/probe_e2e.py, profiles/r02_e2e.md.
  // TAHOE_HOST_SLAB="first,growth%,cap" (waves), read when the handle was created, overrides.
  // Other strategies: slabs of about an eighth of the batch, sized in BYTES (<= 256 MB of rows per slot,
  // whatever the row width).  No slab is larger than the batch.
  const size_t rows32 = (num_rows + 31) / 32 * 32;
  const size_t row_bytes = (size_t)C * sizeof(float);
  size_t slab = ((num_rows + 7) / 8 + 31) / 32 * 32;
  slab = std::max<size_t>(slab, std::min<size_t>(65536, (size_t)(32u << 20) / row_bytes / 32 * 32));
  slab = std::min<size_t>(slab, std::max<size_t>((size_t)(256u << 20) / row_bytes / 32 * 32, 32));
  size_t wave = 0;
  int slab_first = 2, slab_growth = 150, slab_cap = 24;
  if ((strategy == TAHOE_STRATEGY_SHARED_FOREST || strategy == TAHOE_STRATEGY_SPLIT_SHARED_FOREST) &&
      (f->quant || (f->qt && f->sp_use_qt)) && f->trees_per_chunk > 0) {
    const bool paired = f->sparse && C > kStagedMaxCols && f->warps > 8;
    wave = (size_t)f->sm_count * (size_t)(paired ? f->warps / 2 : f->warps) * 32;
    double t[TAHOE_NUM_STRATEGIES];
    select_strategy(f, num_rows, t);
    const double kernel_per_row = t[strategy] / (double)num_rows;
    const double copy_per_row = (double)row_bytes / model::kHostCopyRate;
    if (std::isfinite(kernel_per_row) && kernel_per_row > 0.0) {
      const double g = std::min(3.0, std::max(1.25, 1.15 * kernel_per_row / copy_per_row));
      slab_growth = (int)(g * 100.0 + 0.5);
      slab_cap = (int)std::min(128.0, std::max(24.0, 24.0 * (g / 1.5) * (g / 1.5)));
    }
    if (f->tune.slab_first > 0) {
      slab_first = f->tune.slab_first;
      slab_growth = std::max(f->tune.slab_growth, 100);
      slab_cap = std::max(f->tune.slab_cap, slab_first);
    }
    slab = (size_t)slab_cap * wave;
  }
  slab = std::max<size_t>(32, std::min(slab, rows32));
  // an error leaves copies of earlier slabs in flight INTO h_preds: drain the slot streams before returning
  auto fail = [&](int rc) {
    for (int s = 1; s <= kHostSlots; ++s)
      if (f->ws[s].stream) cudaStreamSynchronize(f->ws[s].stream);
    cudaGetLastError();
    return rc;
  };
#define TAHOE_CUDA_SLAB(call)                                                   \
  do {                                                                          \
    cudaError_t e__ = (call);                                                   \
    if (e__ != cudaSuccess) return fail(::tahoe::cuda_fail(e__, #call));        \
  } while (0)
  int total_launches = 0;
  size_t done = 0;
  double slab_waves = slab_first;
  for (int i = 0; done < num_rows; ++i) {
    Workspace& w = f->ws[1 + i % kHostSlots];
    if (!w.stream) TAHOE_CUDA_SLAB(cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking));
    if (w.cap_slab_rows < slab) {
      TAHOE_CUDA_SLAB(cudaStreamSynchronize(w.stream));
      if (w.slab_data) cudaFree(w.slab_data);
      if (w.slab_preds) cudaFree(w.slab_preds);
      w.slab_data = nullptr; w.slab_preds = nullptr; w.cap_slab_rows = 0;
      TAHOE_CUDA_SLAB(cudaMalloc((void**)&w.slab_data, slab * row_bytes));
      TAHOE_CUDA_SLAB(cudaMalloc((void**)&w.slab_preds, slab * sizeof(float)));
      w.cap_slab_rows = slab;
    }
    size_t this_slab = slab;
    if (wave) {
      this_slab = std::min(slab, (size_t)(slab_waves + 0.5) * wave);
      slab_waves = std::min((double)slab_cap, slab_waves * slab_growth / 100.0);
    }
    const size_t n = std::min(this_slab, num_rows - done);
    TAHOE_CUDA_SLAB(cudaMemcpyAsync(w.slab_data, h_data + done * (size_t)C, n * row_bytes,
                                    cudaMemcpyHostToDevice, w.stream));
    PredictArgs a{w.slab_data, n, w.slab_preds, w.slab_preds, nullptr, w.stream, &w};
    if (!copy_only) {
      int rc = predict_impl(f, strategy, a);
      if (rc) return fail(rc);
      total_launches += f->last_launches;
    }
    TAHOE_CUDA_SLAB(cudaMemcpyAsync(h_preds + done, w.slab_preds, n * sizeof(float), cudaMemcpyDeviceToHost,
                                    w.stream));
    done += n;
  }
  for (int s = 1; s <= kHostSlots; ++s)
    if (f->ws[s].stream) TAHOE_CUDA_SLAB(cudaStreamSynchronize(f->ws[s].stream));
#undef TAHOE_CUDA_SLAB
  f->last_launches = total_launches;
  return TAHOE_OK;
}

int tahoe_predict_host(tahoe_forest_t* f, int strategy, const float* h_data, size_t num_rows,
                       float* h_preds) {
  return predict_host_impl(f, strategy, h_data, num_rows, h_preds, false);
}

int tahoe_host_copy_baseline(tahoe_forest_t* f, int strategy, const float* h_data, size_t num_rows,
                             float* h_scratch) {
  return predict_host_impl(f, strategy, h_data, num_rows, h_scratch, true);
}

int tahoe_strategy_applicable(const tahoe_forest_t* f, int strategy, size_t num_rows) {
  if (!f) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  return applicable(f, strategy, num_rows);
}

int tahoe_select_strategy(const tahoe_forest_t* f, size_t num_rows, double* predicted_seconds) {
  if (!f) { set_error("null argument"); return TAHOE_STRATEGY_DIRECT; }
  return select_strategy(f, num_rows, predicted_seconds);
}

int tahoe_set_profiling(tahoe_forest_t* f, int on) {
  if (!f) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  if (on && f->prof_events.empty()) {
    f->prof_events.resize((size_t)kProfRing * 3);
    for (auto& e : f->prof_events) TAHOE_CUDA(cudaEventCreate(&e));
  }
  f->profile = on != 0;
  f->prof_count = 0;
  return TAHOE_OK;
}

int tahoe_profile_read(tahoe_forest_t* f, double* quantize_ms, double* traverse_ms, int* count) {
  if (!f || !quantize_ms || !traverse_ms || !count) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  double q = 0.0, t = 0.0;
  for (int i = 0; i < f->prof_count; ++i) {
    cudaEvent_t* ev = &f->prof_events[(size_t)i * 3];
    TAHOE_CUDA(cudaEventSynchronize(ev[2]));
    float a = 0.f, b = 0.f;
    TAHOE_CUDA(cudaEventElapsedTime(&a, ev[0], ev[1]));
    TAHOE_CUDA(cudaEventElapsedTime(&b, ev[1], ev[2]));
    q += a; t += b;
  }
  *quantize_ms = q; *traverse_ms = t; *count = f->prof_count;
  return TAHOE_OK;
}

int tahoe_last_launch_count(const tahoe_forest_t* f) { return f ? f->last_launches : 0; }

int tahoe_forest_export(const tahoe_forest_t* f, int which, void* buf, size_t* bytes) {
  if (!f || !bytes) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  const void* src = nullptr;
  size_t n = 0;
  const HostPack& h = f->host;
  switch (which) {
    case TAHOE_EXPORT_HEAP_NODES: src = h.heap.data(); n = h.heap.size() * sizeof(uint2); break;
    case TAHOE_EXPORT_INTERLEAVED: src = h.inter.data(); n = h.inter.size() * sizeof(uint2); break;
    case TAHOE_EXPORT_TREE_ORDER: src = h.tree_order.data(); n = h.tree_order.size() * 4; break;
    case TAHOE_EXPORT_ORIG_IDS: src = h.orig_ids.data(); n = h.orig_ids.size() * 4; break;
    case TAHOE_EXPORT_QUANT_WORDS: src = h.qwords.data(); n = h.qwords.size() * 4; break;
    case TAHOE_EXPORT_QUANT_TABLE: src = h.qtable.data(); n = h.qtable.size() * 4; break;
    case TAHOE_EXPORT_QUANT_OFFSETS: src = h.qoffsets.data(); n = h.qoffsets.size() * 4; break;
    case TAHOE_EXPORT_LEAF_ORIGIN: src = h.leaf_origin.data(); n = h.leaf_origin.size() * 4; break;
    case TAHOE_EXPORT_SIMHASH: src = h.simhash.data(); n = h.simhash.size() * 8; break;
    case TAHOE_EXPORT_QUANT_LUT: src = h.qlut.data(); n = h.qlut.size() * 4; break;
    case TAHOE_EXPORT_QUANT_LUT_MIN: src = h.qlut_min.data(); n = h.qlut_min.size() * 4; break;
    case TAHOE_EXPORT_QUANT_LUT_SCALE: src = h.qlut_scale.data(); n = h.qlut_scale.size() * 4; break;
    case TAHOE_EXPORT_SD_TOP_NODES: src = h.sd_top_nodes.data(); n = h.sd_top_nodes.size() * sizeof(uint2); break;
    case TAHOE_EXPORT_SD_BLOCKS: src = h.sd_blocks.data(); n = h.sd_blocks.size() * sizeof(uint4); break;
    case TAHOE_EXPORT_SD_ORIG_BOTTOM: src = h.sd_orig_bottom.data(); n = h.sd_orig_bottom.size() * 4; break;
    case TAHOE_EXPORT_SPARSE_BFS: src = h.sp_full.data(); n = h.sp_full.size() * sizeof(uint2); break;
    case TAHOE_EXPORT_SPARSE_BFS_ORIG: src = h.sp_orig.data(); n = h.sp_orig.size() * 4; break;
    case TAHOE_EXPORT_SPARSE_TOP: src = h.sp_top.data(); n = h.sp_top.size() * sizeof(uint2); break;
    case TAHOE_EXPORT_SPARSE_DEEP_BLOCKS: src = h.sp_qblocks.data(); n = h.sp_qblocks.size() * sizeof(uint4); break;
    case TAHOE_EXPORT_SPARSE_BLOCK_ORIG: src = h.sp_qblock_orig.data(); n = h.sp_qblock_orig.size() * 4; break;
    case TAHOE_EXPORT_SPARSE_BLOCK_OFFSETS: src = h.sp_qblock_offsets.data(); n = h.sp_qblock_offsets.size() * 8; break;
    default: set_error("unknown export id"); return TAHOE_ERR_INVALID;
  }
  if (buf) {
    if (*bytes < n) { set_error("export buffer too small"); return TAHOE_ERR_INVALID; }
    if (n) std::memcpy(buf, src, n);
  }
  *bytes = n;
  return TAHOE_OK;
}

}  // extern "C"
