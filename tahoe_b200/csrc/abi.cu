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
  if (w.stream) cudaStreamDestroy(w.stream);
  w = Workspace{};
}

static bool whole_forest_fits(const tahoe_forest* f) { return f->quant && f->num_chunks == 1; }

static int applicable(const tahoe_forest* f, int strategy, size_t rows) {
  (void)rows;
  switch (strategy) {
    case TAHOE_STRATEGY_DIRECT:
      return TAHOE_OK;
    case TAHOE_STRATEGY_SHARED_DATA_PER_THREAD:
      return (size_t)f->params.num_cols * 256 * 4 <= (size_t)f->smem_optin ? TAHOE_OK : TAHOE_ERR_NOT_APPLICABLE;
    case TAHOE_STRATEGY_SHARED_DATA:
      return (!f->sparse && f->d_inter && 2 * (((size_t)f->params.num_cols * 4 + 15) & ~(size_t)15) + 208 <= (size_t)f->smem_optin)
                 ? TAHOE_OK : TAHOE_ERR_NOT_APPLICABLE;
    case TAHOE_STRATEGY_SHARED_FOREST:   // whole forest resident (BaseTahoeTest.h:657-668)
      return whole_forest_fits(f) ? TAHOE_OK : TAHOE_ERR_NOT_APPLICABLE;
    case TAHOE_STRATEGY_SPLIT_SHARED_FOREST:  // at least one tree per chunk (:670-679)
      return (f->quant && f->trees_per_chunk >= 1) ? TAHOE_OK : TAHOE_ERR_NOT_APPLICABLE;
    default:
      return TAHOE_ERR_INVALID;
  }
}

// Performance model: per-strategy time = fixed launch cost + node visits / (visit rate at
// full occupancy * fraction of the machine the batch can fill) + row-staging bytes / HBM.
// Same functional form as the reference (T_smem + T_gmem per strategy, main.cu:35-63);
// the rates are B200 measurements (DESIGN.md "Performance model"), not cudaMemcpy guesses.
struct ModelConst {
  double visits_per_s;   // full-machine node visits per second
  double fill_rows;      // rows needed to saturate the machine
  double fixed_s;        // launch + prologue
};
static const ModelConst kModel[TAHOE_NUM_STRATEGIES] = {
    /* DIRECT            */ {4.0e11, 148.0 * 2048, 6e-6},
    /* PER_THREAD        */ {5.0e11, 148.0 * 2048, 8e-6},
    /* SHARED_DATA       */ {2.5e11, 148.0 * 64, 6e-6},
    /* SHARED_FOREST     */ {2.6e12, 148.0 * 512, 12e-6},
    /* SPLIT_SHARED_FOREST*/ {2.6e12, 148.0 * 512, 12e-6},
};

static int select_strategy(const tahoe_forest* f, size_t rows, double* out) {
  const double visits = (double)rows * f->params.num_trees * std::max(f->mean_depth, 1.0);
  const double hbm = 6.5e12;
  int best = TAHOE_STRATEGY_DIRECT;
  double best_t = std::numeric_limits<double>::infinity();
  for (int s = 0; s < TAHOE_NUM_STRATEGIES; ++s) {
    double t = std::numeric_limits<double>::infinity();
    if (applicable(f, s, rows) == TAHOE_OK) {
      const ModelConst& m = kModel[s];
      const double fill = std::min(1.0, std::max((double)rows, 1.0) / m.fill_rows);
      double v = visits;
      if (s >= TAHOE_STRATEGY_SHARED_FOREST) v = (double)rows * f->params.num_trees * f->depth;  // padded
      t = m.fixed_s + v / (m.visits_per_s * fill);
      double bytes = (double)rows * f->params.num_cols * 4.0;
      if (s >= TAHOE_STRATEGY_SHARED_FOREST) {
        bytes = bytes * 2.0 + bytes * f->num_chunks;      // quantise pass + one tile stream per chunk
        t += f->num_chunks * 4e-6;                        // chunk staging + barrier
      }
      t = std::max(t, m.fixed_s + bytes / hbm);
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
  switch (strategy) {
    case TAHOE_STRATEGY_DIRECT: rc = launch_direct(f, a, false, &launches); break;
    case TAHOE_STRATEGY_SHARED_DATA_PER_THREAD: rc = launch_direct(f, a, true, &launches); break;
    case TAHOE_STRATEGY_SHARED_DATA: rc = launch_shared_data(f, a, &launches); break;
    default: rc = launch_smem_forest(f, a, &launches); break;
  }
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

void tahoe_forest_destroy(tahoe_forest_t* f) {
  if (!f) return;
  cudaFree(f->d_heap); cudaFree(f->d_sparse); cudaFree(f->d_tree_offsets); cudaFree(f->d_inter);
  cudaFree(f->d_tree_order); cudaFree(f->d_orig_ids); cudaFree(f->d_qwords); cudaFree(f->d_qtable);
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
  info->trees_per_chunk = f->trees_per_chunk; info->num_chunks = f->num_chunks;
  info->bytes_heap = (int64_t)(f->host.heap.size() * sizeof(uint2) + f->host.sparse.size() * sizeof(uint4));
  info->bytes_interleaved = (int64_t)(f->host.inter.size() * sizeof(uint2));
  info->bytes_quantized = (int64_t)(f->host.qwords.size() * sizeof(uint32_t));
  info->exchanged_nodes = f->exchanged; info->mean_depth = f->mean_depth; info->device = f->device;
  info->sm_count = f->sm_count; info->smem_optin_bytes = f->smem_optin; info->col_shift = f->col_shift;
  info->smem_warps = f->warps; info->smem_staged = f->staged; info->smem_group = f->group;
  return TAHOE_OK;
}

int tahoe_predict(tahoe_forest_t* f, int strategy, const float* d_data, size_t num_rows,
                  float* d_preds, int32_t* d_leaf_ids, void* stream) {
  if (!f || (num_rows && (!d_data || !d_preds))) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  PredictArgs a{d_data, num_rows, d_preds, d_leaf_ids, (cudaStream_t)stream, &f->ws[0]};
  return predict_impl(f, strategy, a);
}

int tahoe_predict_host(tahoe_forest_t* f, int strategy, const float* h_data, size_t num_rows,
                       float* h_preds) {
  if (!f || (num_rows && (!h_data || !h_preds))) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  if (num_rows == 0) return TAHOE_OK;
  const int C = f->params.num_cols;
  if (strategy == TAHOE_STRATEGY_AUTO) strategy = select_strategy(f, num_rows, nullptr);
  size_t slab = ((num_rows + 7) / 8 + 31) / 32 * 32;
  slab = std::min<size_t>(std::max<size_t>(slab, 65536), 1u << 20);
  int total_launches = 0;
  size_t done = 0;
  for (int i = 0; done < num_rows; ++i) {
    Workspace& w = f->ws[1 + i % kHostSlots];
    if (!w.stream) TAHOE_CUDA(cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking));
    if (w.cap_slab_rows < slab) {
      if (w.slab_data) cudaFree(w.slab_data);
      if (w.slab_preds) cudaFree(w.slab_preds);
      w.slab_data = nullptr; w.slab_preds = nullptr; w.cap_slab_rows = 0;
      TAHOE_CUDA(cudaMalloc((void**)&w.slab_data, slab * (size_t)C * sizeof(float)));
      TAHOE_CUDA(cudaMalloc((void**)&w.slab_preds, slab * sizeof(float)));
      w.cap_slab_rows = slab;
    }
    const size_t n = std::min(slab, num_rows - done);
    TAHOE_CUDA(cudaMemcpyAsync(w.slab_data, h_data + done * (size_t)C, n * (size_t)C * sizeof(float),
                               cudaMemcpyHostToDevice, w.stream));
    PredictArgs a{w.slab_data, n, w.slab_preds, nullptr, w.stream, &w};
    int rc = predict_impl(f, strategy, a);
    if (rc) return rc;
    total_launches += f->last_launches;
    TAHOE_CUDA(cudaMemcpyAsync(h_preds + done, w.slab_preds, n * sizeof(float), cudaMemcpyDeviceToHost,
                               w.stream));
    done += n;
  }
  for (int s = 1; s <= kHostSlots; ++s)
    if (f->ws[s].stream) TAHOE_CUDA(cudaStreamSynchronize(f->ws[s].stream));
  f->last_launches = total_launches;
  return TAHOE_OK;
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
