/*
 * tahoe_oracle.c -- CPU restatement of the reference's batch-predict path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this file's shared object, and only as the checker.  The product
 * (tahoe_b200/) never links, imports or calls it.
 *
 * Parity status: PINNED.  Every function below is checked bit-for-bit against
 * the reference's own executable CPU code (oracle/ref_shim.cu compiles
 * /root/reference/BaseTahoeTest.h in place and calls its loaders and
 * infer_one_tree); the resulting vectors are committed under tests/golden/ by
 * tests/golden/make_golden.py.  The reference ships no golden vectors of its
 * own, so its executable CPU path is the pin.
 *
 * Each function cites the reference file:line it restates.  No reference source
 * text is copied; the arithmetic is re-expressed in plain C.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- node encodings --------------------------------------------------------
 * dense_node_t   : Struct.h:44-48  {float weight; float val; int bits;}
 * sparse_node_t  : Struct.h:50-54  {float val; int bits; int left_idx;}
 * bits           : Struct.h:57-59  fid = bits & (2^30-1), def_left = bit 30,
 *                                  is_leaf = bit 31 (encode_node, Struct.h:103-108)
 * output_t flags : Struct.h:37-42  RAW 0, AVG 0x1, SIGMOID 0x10, THRESHOLD 0x100
 */
typedef struct { float weight; float val; int32_t bits; } oracle_dense_node_t;
typedef struct { float val; int32_t bits; int32_t left_idx; } oracle_sparse_node_t;

#define ORACLE_FID_MASK   0x3FFFFFFF
#define ORACLE_DEF_LEFT   0x40000000
#define ORACLE_IS_LEAF    ((int32_t)0x80000000)
#define ORACLE_OUT_AVG        0x1
#define ORACLE_OUT_SIGMOID    0x10
#define ORACLE_OUT_THRESHOLD  0x100

/* Struct.h:15-17: nodes in a complete tree with `depth` internal levels. */
static int64_t tree_nodes(int depth) { return ((int64_t)1 << (depth + 1)) - 1; }

/* Struct.h:13: sigmoid(x) = 1/(1+expf(-x)), all float. */
static float oracle_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }

/* The branch predicate shared by every traversal in the reference
 * (BaseTahoeTest.h:450-452; device copies Struct.h:383, 976-978):
 *   missing test is a FLOAT subtract, fabs, compare against float(1e-6);
 *   otherwise ties go right (>=); a NaN feature fails both tests => left. */
static int branch_right(float val, float thr, int def_left, float missing) {
    const float eps = 1.0e-6f;
    if (fabsf(val - missing) <= eps) return !def_left;
    return val >= thr;
}

/* BaseTahoeTest.h:440-456 infer_one_tree: walk one dense tree for one row.
 * Returns the leaf value; *leaf_id = heap index (root 0, children 2k+1/2k+2)
 * of the first leaf reached -- `curr` at BaseTahoeTest.h:449, which the
 * reference computes but never reports. */
float oracle_infer_one_tree(const oracle_dense_node_t* root, const float* row,
                            float missing, int32_t* leaf_id) {
    int32_t curr = 0;
    for (;;) {
        const oracle_dense_node_t* n = &root[curr];
        if (n->bits & ORACLE_IS_LEAF) {
            if (leaf_id) *leaf_id = curr;
            return n->val;
        }
        int fid = n->bits & ORACLE_FID_MASK;
        int def_left = (n->bits & ORACLE_DEF_LEFT) != 0;
        curr = 2 * curr + 1 + branch_right(row[fid], n->val, def_left, missing);
    }
}

/* Sparse (compact) tree walk: child addressing of Struct.h:2244
 * (curr = left_idx + cond, leaf/def_left bits Struct.h:2231-2233) combined with
 * the eps-sentinel missing test of BaseTahoeTest.h:451-452 (the
 * oracle semantics are the CPU ones, not the isnan() of Struct.h:2240).
 * left_idx is relative to the tree's first node (BaseTahoeTest.h:744). */
float oracle_infer_one_tree_sparse(const oracle_sparse_node_t* root, const float* row,
                                   float missing, int32_t* leaf_id) {
    int32_t curr = 0;
    for (;;) {
        const oracle_sparse_node_t* n = &root[curr];
        if (n->bits & ORACLE_IS_LEAF) {
            if (leaf_id) *leaf_id = curr;
            return n->val;
        }
        int fid = n->bits & ORACLE_FID_MASK;
        int def_left = (n->bits & ORACLE_DEF_LEFT) != 0;
        curr = n->left_idx + branch_right(row[fid], n->val, def_left, missing);
    }
}

/* BaseTahoeTest.h:467-472: AVG (a float DIVISION by num_trees on the CPU path),
 * then +bias, then sigmoid, then threshold -- in that order. */
static float oracle_finish(float pred, int num_trees, int output, float threshold,
                           float global_bias) {
    if (output & ORACLE_OUT_AVG) pred = pred / (float)num_trees;
    pred += global_bias;
    if (output & ORACLE_OUT_SIGMOID) pred = oracle_sigmoid(pred);
    if (output & ORACLE_OUT_THRESHOLD) pred = pred > threshold ? 1.0f : 0.0f;
    return pred;
}

/* BaseTahoeTest.h:458-474 predict_on_cpu: rows outer, trees inner in order
 * 0..T-1, float accumulation starting from 0.0f.  leaf_ids (nullable) is
 * [rows][trees] int32.  Rows are independent, so the optional OpenMP split over
 * rows cannot change any result. */
void oracle_predict_dense(const oracle_dense_node_t* nodes, int num_trees, int depth,
                          const float* data, int64_t num_rows, int num_cols,
                          float missing, int output, float threshold, float global_bias,
                          float* preds, int32_t* leaf_ids) {
    const int64_t npt = tree_nodes(depth);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < num_rows; ++i) {
        float pred = 0.0f;
        for (int j = 0; j < num_trees; ++j) {
            int32_t lid;
            pred += oracle_infer_one_tree(&nodes[(int64_t)j * npt],
                                          &data[i * (int64_t)num_cols], missing, &lid);
            if (leaf_ids) leaf_ids[i * (int64_t)num_trees + j] = lid;
        }
        preds[i] = oracle_finish(pred, num_trees, output, threshold, global_bias);
    }
}

/* Same loop nest over compact trees; tree t starts at nodes[tree_offsets[t]]. */
void oracle_predict_sparse(const oracle_sparse_node_t* nodes, const int64_t* tree_offsets,
                           int num_trees, const float* data, int64_t num_rows,
                           int num_cols, float missing, int output, float threshold,
                           float global_bias, float* preds, int32_t* leaf_ids,
                           int64_t* visits /* nullable: total internal nodes visited */) {
    int64_t total_visits = 0;
#pragma omp parallel for schedule(static) reduction(+ : total_visits)
    for (int64_t i = 0; i < num_rows; ++i) {
        float pred = 0.0f;
        for (int j = 0; j < num_trees; ++j) {
            const oracle_sparse_node_t* root = &nodes[tree_offsets[j]];
            const float* row = &data[i * (int64_t)num_cols];
            int32_t curr = 0;
            for (;;) {
                const oracle_sparse_node_t* n = &root[curr];
                if (n->bits & ORACLE_IS_LEAF) break;
                int fid = n->bits & ORACLE_FID_MASK;
                int def_left = (n->bits & ORACLE_DEF_LEFT) != 0;
                curr = n->left_idx + branch_right(row[fid], n->val, def_left, missing);
                ++total_visits;
            }
            pred += root[curr].val;
            if (leaf_ids) leaf_ids[i * (int64_t)num_trees + j] = curr;
        }
        preds[i] = oracle_finish(pred, num_trees, output, threshold, global_bias);
    }
    if (visits) *visits = total_visits;
}

/* ---- text formats ----------------------------------------------------------
 * BaseTahoeTest.h:267-315 model text: line 1 = num_trees (atoi), line 2 =
 * LEVELS with depth = LEVELS-1 (:281-282), then per tree, per heap node, five
 * lines: fid (atoi), value (atof -> float), default_left (atoi -> bool),
 * weight (atof -> float), is_leaf (atoi -> bool).  Lines are read with
 * fgets into a 1024-byte buffer (:269-270); unparsable text yields 0.
 * Node packing: Struct.h:103-108. */
#define ORACLE_MAX_LINE 1024

int oracle_load_model_text(const char* path, oracle_dense_node_t** nodes_out,
                           int* num_trees_out, int* depth_out) {
    char buf[ORACLE_MAX_LINE];
    FILE* fp = fopen(path, "r");
    if (!fp) return -1;
    int num_trees = 0, depth = 0;
    if (fgets(buf, ORACLE_MAX_LINE, fp)) num_trees = atoi(buf);
    if (fgets(buf, ORACLE_MAX_LINE, fp)) depth = atoi(buf) - 1;
    if (num_trees < 0 || depth < 0 || depth > 30) { fclose(fp); return -2; }
    const int64_t npt = tree_nodes(depth);
    const int64_t total = npt * num_trees;
    oracle_dense_node_t* nodes = (oracle_dense_node_t*)calloc((size_t)(total ? total : 1), sizeof(*nodes));
    if (!nodes) { fclose(fp); return -3; }
    for (int64_t i = 0; i < total; ++i) {
        int fid = 0, def_left = 0, is_leaf = 0;
        float value = 0.0f, weight = 0.0f;
        /* a short file leaves buf unchanged, exactly like the unchecked fgets
         * calls of the reference (:298-307) */
        if (fgets(buf, ORACLE_MAX_LINE, fp)) {} fid = atoi(buf);
        if (fgets(buf, ORACLE_MAX_LINE, fp)) {} value = (float)atof(buf);
        if (fgets(buf, ORACLE_MAX_LINE, fp)) {} def_left = atoi(buf) != 0;
        if (fgets(buf, ORACLE_MAX_LINE, fp)) {} weight = (float)atof(buf);
        if (fgets(buf, ORACLE_MAX_LINE, fp)) {} is_leaf = atoi(buf) != 0;
        nodes[i].weight = weight;
        nodes[i].val = value;
        nodes[i].bits = (fid & ORACLE_FID_MASK) | (def_left ? ORACLE_DEF_LEFT : 0) |
                        (is_leaf ? ORACLE_IS_LEAF : 0);
    }
    fclose(fp);
    *nodes_out = nodes;
    *num_trees_out = num_trees;
    *depth_out = depth;
    return 0;
}

/* BaseTahoeTest.h:354-388 data text: rows (atoi), cols (atoi), missing sentinel
 * (atof -> float), then rows*cols values, row-major, one per line. */
int oracle_load_data_text(const char* path, float** data_out, int* rows_out, int* cols_out,
                          float* missing_out) {
    char buf[ORACLE_MAX_LINE];
    FILE* fp = fopen(path, "r");
    if (!fp) return -1;
    int rows = 0, cols = 0;
    float missing = 0.0f;
    if (fgets(buf, ORACLE_MAX_LINE, fp)) rows = atoi(buf);
    if (fgets(buf, ORACLE_MAX_LINE, fp)) cols = atoi(buf);
    if (fgets(buf, ORACLE_MAX_LINE, fp)) missing = (float)atof(buf);
    if (rows < 0 || cols < 0) { fclose(fp); return -2; }
    const int64_t total = (int64_t)rows * cols;
    float* data = (float*)calloc((size_t)(total ? total : 1), sizeof(float));
    if (!data) { fclose(fp); return -3; }
    for (int64_t i = 0; i < total; ++i) {
        if (fgets(buf, ORACLE_MAX_LINE, fp)) {}
        data[i] = (float)atof(buf);
    }
    fclose(fp);
    *data_out = data;
    *rows_out = rows;
    *cols_out = cols;
    *missing_out = missing;
    return 0;
}

void oracle_free(void* p) { free(p); }

/* BaseTahoeTest.h:727-757 (commented-out dense2sparse*): depth-first conversion
 * of one dense tree into compact nodes, children allocated as an adjacent pair,
 * left_idx relative to the tree root.  Used by tests to derive sparse fixtures
 * from dense ones.  Returns the number of nodes written (<= cap) or -1. */
static int64_t d2s_node(const oracle_dense_node_t* dense, int64_t i_dense,
                        oracle_sparse_node_t* out, int64_t* used, int64_t cap,
                        int64_t i_sparse) {
    const oracle_dense_node_t* n = &dense[i_dense];
    out[i_sparse].val = n->val;
    out[i_sparse].bits = n->bits;
    out[i_sparse].left_idx = 0;
    if (n->bits & ORACLE_IS_LEAF) return 0;
    if (*used + 2 > cap) return -1;
    int64_t left = *used;
    *used += 2;
    out[i_sparse].left_idx = (int32_t)left;
    if (d2s_node(dense, 2 * i_dense + 1, out, used, cap, left) < 0) return -1;
    if (d2s_node(dense, 2 * i_dense + 2, out, used, cap, left + 1) < 0) return -1;
    return 0;
}

int64_t oracle_dense_to_sparse_tree(const oracle_dense_node_t* dense_root,
                                    oracle_sparse_node_t* out, int64_t cap) {
    if (cap < 1) return -1;
    int64_t used = 1;
    if (d2s_node(dense_root, 0, out, &used, cap, 0) < 0) return -1;
    return used;
}
