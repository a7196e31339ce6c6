// ref_shim.cu -- thin C entry points around the REFERENCE's own host code.
//
// TEST INFRASTRUCTURE ONLY (see oracle/tahoe_oracle.c header).  This file holds
// no algorithm: it #includes /root/reference/BaseTahoeTest.h where it lies
// (read-only, never copied into the repo) and calls the reference's public
// members.  It is built by oracle/Makefile into oracle/_ref/libtahoe_ref.so,
// which is git-ignored and travels to the GPU box as a prebuilt file.
//
// What runs here is the reference's executable CPU path:
//   generate_forest_from_file   BaseTahoeTest.h:267-352
//   generate_data_from_file     BaseTahoeTest.h:354-402
//   infer_one_tree              BaseTahoeTest.h:440-456
//   dense_adaptive_forest::init Struct.h:1756-1986   (packer, host arrays only)
// predict_on_cpu (BaseTahoeTest.h:458-487) keeps its result in a local vector
// and only uploads it, so ref_predict() re-states its row/tree double loop
// (:462-472) around the reference's own infer_one_tree -- the arithmetic that
// decides every branch and every leaf value is the reference's.
//
// Without a GPU the CUDA calls inside the reference fail, print "FAIL: ..." and
// continue (cuda_base.h:19-25); the host arrays stay valid.
#include "BaseTahoeTest.h"

#include <cstdint>
#include <cstring>
#include <unistd.h>

namespace {
BaseTahoeTest* make_harness(const char* model, const char* data) {
  return new BaseTahoeTest(std::string(model ? model : ""), std::string(data ? data : ""));
}
}  // namespace

extern "C" {

int ref_sizeof_dense_node() { return (int)sizeof(dense_node_t); }

// Parse a model text file with the reference's loader.  Returns 0 on success.
int ref_load_model(const char* path, dense_node_t** nodes_out, int* num_trees, int* depth) {
  if (access(path, R_OK) != 0) return -1;  // the reference would exit(1)
  BaseTahoeTest* t = make_harness(path, "");
  t->generate_forest_from_file();
  size_t n = t->nodes.size();
  dense_node_t* out = (dense_node_t*)malloc(sizeof(dense_node_t) * (n ? n : 1));
  memcpy(out, t->nodes.data(), sizeof(dense_node_t) * n);
  *nodes_out = out;
  *num_trees = t->ps.num_trees;
  *depth = t->ps.depth;
  delete t;
  return 0;
}

// Parse a data text file with the reference's loader (its device upload fails
// harmlessly on a GPU-less host).
int ref_load_data(const char* path, float** data_out, int* rows, int* cols, float* missing) {
  if (access(path, R_OK) != 0) return -1;
  BaseTahoeTest* t = make_harness("", path);
  t->stream = 0;
  t->generate_data_from_file();
  size_t n = t->data_h.size();
  float* out = (float*)malloc(sizeof(float) * (n ? n : 1));
  memcpy(out, t->data_h.data(), sizeof(float) * n);
  *data_out = out;
  *rows = t->ps.num_rows;
  *cols = t->ps.num_cols;
  *missing = t->ps.missing;
  if (t->data_d) cudaFree(t->data_d);
  delete t;
  return 0;
}

void ref_free(void* p) { free(p); }

// The reference's "standard results": BaseTahoeTest.h:462-472 around its own
// infer_one_tree.  nthreads <= 1 is the reference as shipped (single thread);
// nthreads > 1 splits the independent rows over OpenMP threads.
void ref_predict(const dense_node_t* nodes, int num_trees, int depth, const float* data,
                 long num_rows, int num_cols, float missing, int output, float threshold,
                 float global_bias, float* preds, int nthreads) {
  BaseTahoeTest* t = make_harness("", "");
  t->ps.num_trees = num_trees;
  t->ps.depth = depth;
  t->ps.num_cols = num_cols;
  t->ps.missing = missing;
  t->ps.output = (output_t)output;
  t->ps.threshold = threshold;
  t->ps.global_bias = global_bias;
  const long npt = t->tree_num_nodes();
  dense_node_t* nd = const_cast<dense_node_t*>(nodes);
  float* dt = const_cast<float*>(data);
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(static)
  for (long i = 0; i < num_rows; ++i) {
    float pred = 0.0f;
    for (int j = 0; j < num_trees; ++j) {
      pred += t->infer_one_tree(&nd[(long)j * npt], &dt[i * (long)num_cols]);
    }
    if ((t->ps.output & output_t::AVG) != 0) pred = pred / t->ps.num_trees;
    pred += t->ps.global_bias;
    if ((t->ps.output & output_t::SIGMOID) != 0) pred = sigmoid(pred);
    if ((t->ps.output & output_t::THRESHOLD) != 0) pred = pred > t->ps.threshold ? 1.0f : 0.0f;
    preds[i] = pred;
  }
  delete t;
}

// Run the reference's packer (Struct.h:1756-1986) and hand back its host-side
// tree-major SoA arrays (bits_values_h / bits_{char,short,int}_h), i.e. the
// nodes after the probability reorder with the exchange bit set, in the
// reference's (undefined) tree order.  bits are widened to int32.
// Returns bits_length (1, 2, 4) or -1.
int ref_pack(const dense_node_t* nodes, int num_trees, int depth, int num_cols,
             float* values_out, int32_t* bits_out) {
  forest_params_t p;
  memset(&p, 0, sizeof(p));
  p.depth = depth;
  p.num_trees = num_trees;
  p.num_cols = num_cols;
  p.algo = algo_t::NAIVE;
  p.output = output_t::RAW;
  p.strategy = strategy_t::SHARED_DATA;
  dense_adaptive_forest* f = new dense_adaptive_forest;
  f->init(0, nodes, &p);
  const long n = (long)num_trees * (((long)1 << (depth + 1)) - 1);
  int bl = f->bits_length;
  for (long i = 0; i < n; ++i) {
    values_out[i] = f->bits_values_h[i];
    if (bl == 1) bits_out[i] = (int32_t)(unsigned char)f->bits_char_h[i];
    else if (bl == 2) bits_out[i] = (int32_t)(unsigned short)f->bits_short_h[i];
    else if (bl == 4) bits_out[i] = f->bits_int_h[i];
    else bits_out[i] = 0;
  }
  return bl;
}

}  // extern "C"
