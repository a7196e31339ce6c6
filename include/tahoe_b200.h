/*
 * tahoe_b200.h -- C ABI of the B200-native decision-forest batch-predict engine.
 *
 * Drop-in boundary for ONE path of sampathrg/Tahoe: forest load -> traversal
 * strategies -> per-row summed score.  The reference has no FFI layer; its
 * boundary is a header-only C++ class plus an abstract struct.
 * Every entry point below names the reference interface it replaces.
 *
 * Plain pointers and sizes only; no torch / C++ types.  All functions return a
 * tahoe_status_t (0 = ok) unless stated otherwise and never abort; the text of
 * the last failure on the calling thread is available from tahoe_last_error().
 * Every compute entry point runs on the GPU; there is no CPU fallback.
 */
#ifndef TAHOE_B200_H
#define TAHOE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes ------------------------------------------------------------
 * The reference prints and continues (cuda_base.h:19-25) or assert(false)s
 * (Struct.h:2012, 2057, 2114); here everything is a status. */
typedef enum {
  TAHOE_OK = 0,
  TAHOE_ERR_INVALID = 1,        /* bad argument / malformed forest */
  TAHOE_ERR_CUDA = 2,           /* CUDA runtime failure (text in tahoe_last_error) */
  TAHOE_ERR_NOT_APPLICABLE = 3, /* strategy cannot run this forest/batch:
                                   "Strategy %d is not suitable for this case."
                                   (BaseTahoeTest.h:662-678) */
  TAHOE_ERR_IO = 4,             /* unreadable file (reference: exit(1), BaseTahoeTest.h:273-277) */
  TAHOE_ERR_NOMEM = 5
} tahoe_status_t;

/* ---- strategies (value = reference `selected_algorithm`,
 * Struct.h:2168-2179; SetUp()'s printed id is value+1, BaseTahoeTest.h:93-109) */
typedef enum {
  TAHOE_STRATEGY_AUTO = -1,                  /* performance model picks (main.cu:22-82) */
  TAHOE_STRATEGY_DIRECT = 0,                 /* K1  Struct.h:1196-1240 */
  TAHOE_STRATEGY_SHARED_DATA_PER_THREAD = 1, /* K2  Struct.h:1610-1704 */
  TAHOE_STRATEGY_SHARED_DATA = 2,            /* K3  Struct.h:992-1191  */
  TAHOE_STRATEGY_SHARED_FOREST = 3,          /* K4  Struct.h:1245-1409 */
  TAHOE_STRATEGY_SPLIT_SHARED_FOREST = 4,    /* K5  Struct.h:1414-1606 */
  TAHOE_NUM_STRATEGIES = 5
} tahoe_strategy_t;

/* output_t flags, Struct.h:37-42 */
enum { TAHOE_OUT_RAW = 0x0, TAHOE_OUT_AVG = 0x1, TAHOE_OUT_SIGMOID = 0x10, TAHOE_OUT_THRESHOLD = 0x100 };

/* ---- POD layouts shared with the reference ------------------------------------
 * dense_node_t, Struct.h:44-48.  bits = fid[0:29] | def_left<<30 | is_leaf<<31
 * (Struct.h:57-59, encode_node :103-108).  Trees are complete heaps of
 * 2^(depth+1)-1 nodes (root 0, children 2k+1 / 2k+2), tree-major. */
typedef struct { float weight; float val; int32_t bits; } tahoe_dense_node_t;
/* sparse_node_t, Struct.h:50-54.  Children are an adjacent pair at
 * left_idx / left_idx+1 relative to the tree's first node (Struct.h:2244). */
typedef struct { float val; int32_t bits; int32_t left_idx; } tahoe_sparse_node_t;

/* forest_params_t, Struct.h:166-189 (same members, same order; enums as int). */
typedef struct {
  int num_nodes;     /* total nodes; sparse forests only */
  int depth;         /* internal levels; dense forests only */
  int num_trees;
  int num_cols;
  int algo;          /* algo_t, accepted and ignored (layouts are chosen per strategy) */
  int output;        /* TAHOE_OUT_* flags */
  float threshold;
  float global_bias;
  int strategy;      /* strategy_t of the FIL baseline, accepted and ignored */
  float missing;     /* missing-value sentinel (data file header line 3) */
} tahoe_forest_params_t;

typedef struct tahoe_forest tahoe_forest_t; /* opaque handle; owns all device buffers */

typedef struct {
  int num_trees, depth, num_cols, is_sparse;
  int64_t num_nodes;          /* real nodes */
  int node_width_bytes;       /* reference width rule b in {1,2,4} (Struct.h:1831-1852), 0 if undefined */
  int quantized;              /* 1 if the rank-quantised 4-byte layout exists */
  int rank_bits, col_bits;
  int trees_per_chunk;        /* shared-memory forest chunk (0 = a tree does not fit) */
  int num_chunks;
  int64_t bytes_heap;         /* packed bytes: 8-byte heap layout (DIRECT, PER_THREAD) */
  int64_t bytes_interleaved;  /* node-major [node][tree] layout (SHARED_DATA) */
  int64_t bytes_quantized;    /* 4-byte rank layout (SHARED_FOREST / SPLIT) */
  int64_t exchanged_nodes;    /* nodes whose children were swapped by the probability reorder */
  double mean_depth;          /* mean internal nodes on a root-to-leaf path, cover-weighted */
  int device;
  int sm_count;
  int smem_optin_bytes;
  int col_shift;              /* bit position of the column field in a quantised node word (7 = compact, else wide) */
  int smem_warps;             /* shared-memory-forest plan: warps per CTA, */
  int smem_staged;            /*   tile buffers per warp: 0 = two by TMA, 1 = one + register staging, 2 = one by TMA, refilled from L2 after the walk, */
  int smem_group;             /*   trees traversed together by one thread */
  int sd_teams;               /* SHARED_DATA plan: row teams per CTA (0 = two rows do not fit shared memory), */
  int sd_team_warps;          /*   warps sharing one row, */
  int sd_top_levels;          /*   forest levels copied to shared memory, */
  int sd_blocked;             /*   1 = the levels below are walked two at a time through 32-byte blocks */
  int sparse_cut;             /* sparse forests through shared memory: breadth-first nodes of each tree kept there */
  double sparse_deep_walks;   /*   expected fraction of (row, tree) walks that go deeper (finished in L2) */
  int sparse_fid_bits;        /*   width of the feature-id field of an 8-byte sparse node (0 = fields do not fit) */
  int sparse_top_depth;       /* sparse forests, quantised tops (QT): levels of every tree kept as a dense 4-byte heap (0 = not built) */
  int sparse_top_active;      /*   1 = SHARED_FOREST / SPLIT_SHARED_FOREST run the QT path (else the 8-byte breadth-first tops) */
  int sparse_top_trees_per_chunk;
  int sparse_top_chunks;
  double sparse_top_deep_walks; /* expected fraction of (row, tree) walks still on an internal node below the top (finished in L2) */
  int smem_cluster;           /* shared-memory-forest plan: CTAs per thread-block cluster (0 = no cluster pipeline) */
} tahoe_forest_info_t;

/* ---- forest lifetime -------------------------------------------------------------
 * Replaces dense_adaptive_forest::init(stream, nodes, params), Struct.h:1756, and
 * init_dense_adaptive, BaseTahoeTest.h:605-611.  `nodes` is the host AoS array of
 * num_trees * (2^(depth+1)-1) entries; it is copied.  The forest is built on the
 * current CUDA device. */
int tahoe_forest_create_dense(const tahoe_dense_node_t* nodes, const tahoe_forest_params_t* params,
                              tahoe_forest_t** out);
/* Replaces sparse_forest::init, Struct.h:2327-2345 (disabled in the reference).
 * tree_offsets has num_trees+1 entries; weights (nullable) are per-node cover
 * fractions used by the probability reorder. */
int tahoe_forest_create_sparse(const tahoe_sparse_node_t* nodes, const float* weights,
                               const int64_t* tree_offsets, const tahoe_forest_params_t* params,
                               tahoe_forest_t** out);
/* Per-class outputs.  The reference sums every tree into ONE scalar, also for multi-class
 * models (Struct.h:196-209, BaseTahoeTest.h:462-466); here tree t belongs to class tree_class[t] in
 * [0, num_classes) (XGBoost's `tree_info`), all classes share one packed forest -- one set of rank tables, so
 * ONE quantise pass per batch serves every class -- and tahoe_predict_classes writes d_out[k][row] = the score of
 * class k: its trees summed in their given order, then the handle's output transform (AVG divides by the class's
 * own tree count).  d_out: device, [num_classes][num_rows] float32.  Strategies: DIRECT,
 * SHARED_DATA_PER_THREAD, SHARED_FOREST / SPLIT_SHARED_FOREST (AUTO: the model's pick; others fall back to DIRECT).
 * tahoe_predict on such a handle still returns the sum over all classes. */
int tahoe_forest_create_sparse_classes(const tahoe_sparse_node_t* nodes, const float* weights,
                                       const int64_t* tree_offsets, const tahoe_forest_params_t* params,
                                       int num_classes, const int32_t* tree_class, tahoe_forest_t** out);
int tahoe_predict_classes(tahoe_forest_t* f, int strategy, const float* d_data, size_t num_rows, float* d_out,
                          void* stream);
/* The reference leaks its forest (`delete forest`, BaseTahoeTest.h:709, no destructor). */
void tahoe_forest_destroy(tahoe_forest_t* f);
int tahoe_forest_info(const tahoe_forest_t* f, tahoe_forest_info_t* info);

/* ---- predict -----------------------------------------------------------------------
 * Replaces forest::predict(stream, preds, data, num_rows), Struct.h:245-269, with the
 * strategy passed explicitly instead of through the global `selected_algorithm`
 * (Struct.h:11).  d_data: device, row-major [num_rows][num_cols] float32.
 * d_preds: device [num_rows].  d_leaf_ids: nullable device [num_rows][num_trees]
 * int32 receiving the heap index (dense) / node index within the tree (sparse) of
 * the leaf each row reaches in each tree, in the caller's original tree order.
 * stream: a cudaStream_t (NULL = default stream).  Asynchronous. */
int tahoe_predict(tahoe_forest_t* f, int strategy, const float* d_data, size_t num_rows,
                  float* d_preds, int32_t* d_leaf_ids, void* stream);
/* tahoe_predict with the FINAL scores going to d_out, which the kernels only ever write (once per
 * row, in the epilogue of the last forest chunk) and never read back: d_out may be memory of
 * another GPU mapped into this process (cudaIpcOpenMemHandle / peer access).  This is the gather of
 * the multi-GPU path fused into the kernel epilogue: rank g passes rank 0's gathered
 * buffer + its own row offset, and no copy or collective follows the predict.  The reference has no
 * counterpart (one stream, one device, BaseTahoeTest.h:78).  d_partial: local scratch of num_rows
 * floats for the partial sums the multi-pass strategies carry between chunks; NULL = scratch owned
 * by the handle (grow-only).
 * Calls on one handle are stream-ordered; calls issued on different streams are serialised on the
 * device by an event (they share the handle's scratch); a handle must not be used from two host
 * threads at once. */
int tahoe_predict_to(tahoe_forest_t* f, int strategy, const float* d_data, size_t num_rows,
                     float* d_out, float* d_partial, int32_t* d_leaf_ids, void* stream);
/* The gathered output buffer of the multi-GPU path: the owner (rank 0) creates it on its device and
 * gets a CUDA IPC handle (TAHOE_IPC_HANDLE_BYTES opaque bytes) to pass to the other processes by any means;
 * every other process opens it WITH ITS OWN DEVICE CURRENT and gets a pointer its kernels can store through
 * (peer mapping over NVLink / NVSwitch), to be used as d_out of tahoe_predict_to at its row offset.
 * release: owner = 1 frees the buffer, owner = 0 closes the mapping.  No reference counterpart (single device,
 * BaseTahoeTest.h:78). */
#define TAHOE_IPC_HANDLE_BYTES 64
int tahoe_gather_buffer_create(size_t num_floats, float** d_ptr, unsigned char* ipc_handle);
int tahoe_gather_buffer_open(const unsigned char* ipc_handle, float** d_ptr);
int tahoe_gather_buffer_release(float* d_ptr, int owner);
/* Lets kernels launched on the CURRENT device store into memory of `peer_device` (the device that
 * owns the gathered output buffer of tahoe_predict_to).  Idempotent.  TAHOE_ERR_NOT_APPLICABLE when
 * the two devices have no peer path. */
int tahoe_enable_peer_access(int peer_device);
/* No counterpart in the reference.
 * The shared-memory-forest predict (k_quantize + k_smem_forest, i.e. tahoe_predict with
 * TAHOE_STRATEGY_SPLIT_SHARED_FOREST) for callers that cannot allocate: code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc.  It calls no cudaMalloc / cudaFree / allocating
 * library routine: the kernels' scratch (the batch's rank tiles and per-tile flag words) is a file-scope
 * static __device__ array of TAHOE_NO_ALLOC_SCRATCH_BYTES that every GPU loading the library holds until
 * exit.  Limit: roundup(4 * tiles, 256) + tiles * 128 * num_cols <= TAHOE_NO_ALLOC_SCRATCH_BYTES with
 * tiles = ceil(num_rows / 32) (232,608 rows of 18 columns, 149,600 of 28, 83,808 of 50); more rows -> TAHOE_ERR_INVALID with the limit in
 * tahoe_last_error().  Forests without the rank-quantised layout -> TAHOE_ERR_NOT_APPLICABLE.  The handle's
 * device must be current.  d_preds is both the partial-sum carrier and the output (local memory).  Same
 * results as tahoe_predict, bit for bit.  ALL calls on one GPU share the array, which keeps the previous
 * call's contents (what must start at zero is cleared on `stream` first): calls that overlap on the GPU --
 * two streams, two handles, two host threads -- race. */
#define TAHOE_NO_ALLOC_SCRATCH_BYTES (16u << 20)
int tahoe_predict_no_alloc(tahoe_forest_t* f, const float* d_data, size_t num_rows, float* d_preds,
                           void* stream);
/* Same call with HOST buffers: slabs of rows are copied in, predicted and copied
 * out on internal streams so that PCIe transfers overlap the traversal.  This is
 * the harness path generate_data_from_file -> updateDevice -> predict -> updateHost
 * (BaseTahoeTest.h:392, 691-692).  Synchronous. */
int tahoe_predict_host(tahoe_forest_t* f, int strategy, const float* h_data, size_t num_rows,
                       float* h_preds);
/* Measurement aid: the slab schedule of tahoe_predict_host for `strategy` with the kernels left out --
 * the same host-to-device and device-to-host copies on the same streams and nothing else -- so a
 * caller can time what the copies alone cost (the bound of the end-to-end path).  h_scratch
 * receives num_rows floats of unspecified content. */
int tahoe_host_copy_baseline(tahoe_forest_t* f, int strategy, const float* h_data, size_t num_rows,
                             float* h_scratch);

/* TAHOE_OK if `strategy` can run this forest on `num_rows` rows, else
 * TAHOE_ERR_NOT_APPLICABLE (harness gates, BaseTahoeTest.h:657-679). */
int tahoe_strategy_applicable(const tahoe_forest_t* f, int strategy, size_t num_rows);
/* Performance model (replaces main.cu:22-82): returns the strategy id it picks for
 * this forest and batch size, and optionally the predicted seconds per strategy
 * (array of TAHOE_NUM_STRATEGIES, +inf where not applicable). */
int tahoe_select_strategy(const tahoe_forest_t* f, size_t num_rows, double* predicted_seconds);
/* Number of kernel launches the last tahoe_predict on this handle issued. */
int tahoe_last_launch_count(const tahoe_forest_t* f);

/* Per-kernel timing for the roofline report.  While on, every tahoe_predict on the
 * shared-memory-forest strategies records CUDA events on the launch stream around its two
 * kernels (the reference times whole predict calls with gettimeofday, BaseTahoeTest.h:687-698).
 * tahoe_profile_read synchronises on them and returns the summed milliseconds and the number
 * of predicts covered since profiling was switched on. */
int tahoe_set_profiling(tahoe_forest_t* f, int on);
int tahoe_profile_read(tahoe_forest_t* f, double* quantize_ms, double* traverse_ms, int* count);

/* ---- roofline denominators --------------------------------------------------------------
 * Read-bandwidth microbenchmarks on the current device (CUDA events, best of 5): conflict-free
 * shared-memory LDS.32 stream, random shared-memory word loads, L2-resident streaming read (32 MB
 * buffer, 16-byte loads) and L2-resident 32-byte sector gathers.  The reference's counterpart is
 * its copy of the CUDA sample bandwidthTest (bandwidthTest.h:110 runTest, main.cu:8-15). */
typedef struct {
  double smem_read_gbs;
  double smem_random_word_loads_per_s;
  double l2_read_gbs;
  double l2_sector_gathers_per_s;
  int sm_count;
  int l2_bytes;
  int persisting_l2_max_bytes;
} tahoe_peaks_t;
int tahoe_measure_peaks(tahoe_peaks_t* out);

/* ---- text loaders (host only) -------------------------------------------------------
 * generate_forest_from_file, BaseTahoeTest.h:267-352 / generate_data_from_file,
 * :354-402 (parse only; the caller uploads).  Buffers are released with
 * tahoe_free_host. */
int tahoe_load_model_text(const char* path, tahoe_dense_node_t** nodes, int* num_trees, int* depth);
int tahoe_load_data_text(const char* path, float** data, int* num_rows, int* num_cols, float* missing);
void tahoe_free_host(void* p);
/* Both loaders map the file and parse it on all host threads (TAHOE_LOADER_THREADS), with the
 * reference's exact per-line semantics ((float)atof / atoi of a 1024-byte fgets buffer, a short
 * file repeating its last line); TAHOE_LOADER_SEQUENTIAL=1 forces the reference's one-thread loop.
 *
 * Binary cache of a parsed file (the reference re-parses its text on every run): a
 * 64-byte header {magic, version, kind, shape, missing, payload bytes, FNV-1a 64 checksum} followed
 * by the raw payload (12-byte AoS nodes / row-major float32).  Loading verifies magic, version,
 * shape against payload size, and the checksum; buffers are released with tahoe_free_host. */
int tahoe_save_model_binary(const char* path, const tahoe_dense_node_t* nodes, int num_trees, int depth);
int tahoe_load_model_binary(const char* path, tahoe_dense_node_t** nodes, int* num_trees, int* depth);
int tahoe_save_data_binary(const char* path, const float* data, int num_rows, int num_cols, float missing);
int tahoe_load_data_binary(const char* path, float** data, int* num_rows, int* num_cols, float* missing);

/* ---- synthetic inputs (device) -------------------------------------------------------
 * Seeded U(-1,1) sample matrix written straight into device memory; bit-identical
 * to tahoe_b200.synth.gen_data (replaces generate_data, BaseTahoeTest.h:404-438). */
int tahoe_gen_data_device(uint64_t seed, size_t row0, size_t num_rows, int num_cols,
                          float miss_frac, float missing, float* d_out, void* stream);

/* ---- introspection for tests -----------------------------------------------------------
 * Copies a packed host-side array out of the handle.  which: see TAHOE_EXPORT_*.
 * Call with buf == NULL to get the size in bytes through *bytes. */
enum {
  TAHOE_EXPORT_HEAP_NODES = 0,    /* uint2 [T][2^(depth+1)] 1-based heap: {val bits, meta} */
  TAHOE_EXPORT_INTERLEAVED = 1,   /* uint2 [nodes][Tpad] after reorder + tree permutation */
  TAHOE_EXPORT_TREE_ORDER = 2,    /* int32 [Tpad] packed slot -> original tree (-1 = padding) */
  TAHOE_EXPORT_ORIG_IDS = 3,      /* int32 [nodes][Tpad] packed position -> original heap id */
  TAHOE_EXPORT_QUANT_WORDS = 4,   /* uint32 [T][2^(depth+1)] rank-quantised trees */
  TAHOE_EXPORT_QUANT_TABLE = 5,   /* float [sum n_col] sorted distinct thresholds per column */
  TAHOE_EXPORT_QUANT_OFFSETS = 6, /* int32 [num_cols+1] offsets into the table */
  TAHOE_EXPORT_LEAF_ORIGIN = 7,   /* int32 [T][2^depth] padded bottom slot -> original leaf heap id */
  TAHOE_EXPORT_SIMHASH = 8,       /* uint64 [T] simhash per original tree */
  TAHOE_EXPORT_QUANT_LUT = 9,     /* int32 [C][B+1] bucket index: thresholds of the column in lower buckets */
  TAHOE_EXPORT_QUANT_LUT_MIN = 10,  /* float [C] */
  TAHOE_EXPORT_QUANT_LUT_SCALE = 11, /* float [C] */
  TAHOE_EXPORT_SD_TOP_NODES = 12,  /* uint2 [2^sd_top_levels - 1][Tpad] leaves pushed down (empty unless sd_blocked) */
  TAHOE_EXPORT_SD_BLOCKS = 13,     /* uint32 [block][Tpad][8] = {root, left, right, 0} x {val bits, word}; word (also in SD_TOP_NODES) = byte offset of the feature in the row | bit 31: a missing value takes child slot 1 |
                                      bit 30: children exchanged (v >= thr takes slot 0) | bit 29: pushed-down copy of a leaf */
  TAHOE_EXPORT_SD_ORIG_BOTTOM = 14, /* int32 [2^depth][Tpad] bottom position -> original heap id of the leaf */
  TAHOE_EXPORT_SPARSE_BFS = 15,      /* uint2 [nodes] sparse trees renumbered breadth-first: {val, leaf<<31|def_left<<30|delta<<fid_bits|fid} */
  TAHOE_EXPORT_SPARSE_BFS_ORIG = 16, /* int32 [nodes] breadth-first index -> caller's node index inside its tree */
  TAHOE_EXPORT_SPARSE_TOP = 17,      /* uint2 [chunks * trees_per_chunk][sparse_cut] shared-memory image of the tree tops */
  TAHOE_EXPORT_SPARSE_DEEP_BLOCKS = 18,  /* uint32 [blocks + 1][8] QT layout below the dense tops: two-level blocks {root, left, right, -} x
                                            {rank word | leaf value, leaf<<31 | block delta}; the last block is the dummy leaf (+0.0f);
                                            the dense tops themselves are exported as QUANT_WORDS / LEAF_ORIGIN */
  TAHOE_EXPORT_SPARSE_BLOCK_ORIG = 19,   /* int32 [blocks + 1][3] caller's node index (inside its tree) of root / left / right, -1 = none */
  TAHOE_EXPORT_SPARSE_BLOCK_OFFSETS = 20 /* int64 [T+1] first block of every tree */
};
int tahoe_forest_export(const tahoe_forest_t* f, int which, void* buf, size_t* bytes);

const char* tahoe_last_error(void);
const char* tahoe_version(void);

#ifdef __cplusplus
}
#endif
#endif /* TAHOE_B200_H */
