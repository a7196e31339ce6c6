// kernels_direct.cu -- thread = sample strategies with the forest read through L1/L2.
//
//   DIRECT                  (reference K1, Struct.h:1196-1240): features straight from global
//   SHARED_DATA_PER_THREAD  (reference K2, Struct.h:1610-1704): the block's rows staged once in
//                           shared memory, COLUMN-major ([col][thread]) so that lane i always
//                           reads bank i -- no bank conflicts whatever feature each lane needs
//                           (the reference copies each row serially, uncoalesced, row-major).
// Both sum leaf values in tree order 0..T-1 exactly like predict_on_cpu
// (BaseTahoeTest.h:462-466), so their raw scores are bit-identical to the CPU's.
#include <cstdlib>

#include "device_utils.cuh"
#include "forest.h"

namespace tahoe {

constexpr int kDirectThreads = 256;

struct DirectParams {
  const float* data;
  size_t rows;
  int cols;
  int num_trees;
  float missing;
  const uint2* heap;            // dense: [T][2^(d+1)]
  int64_t tree_stride;          // 2^(d+1)
  const uint4* sparse;          // sparse nodes, S16 {val, bits, left_idx, 0}
  const uint2* sparse8;         // sparse nodes, S8 {val, leaf|def_left|delta|fid} (when the fields fit)
  int fid_bits;                 // S8: width of the fid field; delta = left_idx - own index sits above it
  const int64_t* tree_offsets;  // sparse
  float* preds;
  int32_t* leaf_ids;
  Transform tf;
};

template <bool SPARSE, bool SMEM_ROWS, int kDirectGroup>
__global__ void __launch_bounds__(kDirectThreads) k_direct(DirectParams p) {
  extern __shared__ float s_rows[];  // SMEM_ROWS: [cols][kDirectThreads]
  const size_t block_row0 = (size_t)blockIdx.x * kDirectThreads;
  const size_t row = block_row0 + threadIdx.x;
  if (SMEM_ROWS) {
    // the block's rows are one contiguous run of floats: coalesced read, transposed store
    const size_t nrows = min((size_t)kDirectThreads, p.rows - block_row0);
    const size_t total = nrows * (size_t)p.cols;
    const float* src = p.data + block_row0 * (size_t)p.cols;
    for (size_t w = threadIdx.x; w < total; w += kDirectThreads) {
      const int r = (int)(w / p.cols), c = (int)(w % p.cols);
      s_rows[c * kDirectThreads + r] = __ldg(src + w);
    }
    __syncthreads();
  }
  if (row >= p.rows) return;
  const float* grow = p.data + row * (size_t)p.cols;
  float acc = 0.0f;
  if (kDirectGroup == 1) {
    // Full machine: plenty of warps hide the L2 latency, and a per-lane loop lets lanes whose walk
    // ended early drop out instead of re-reading their leaf.
    for (int t = 0; t < p.num_trees; ++t) {
      float leaf_val;
      int32_t leaf_id;
      if (!SPARSE) {
        const uint2* tree = p.heap + (int64_t)t * p.tree_stride;
        uint32_t n = 1;
        uint2 e = __ldg(tree + 1);
        while (!(e.y & kMetaLeaf)) {
          const int fid = e.y & 0x3FFFFFFF;
          const float v = SMEM_ROWS ? s_rows[fid * kDirectThreads + threadIdx.x] : __ldg(grow + fid);
          n = 2 * n + (branch_right(v, __uint_as_float(e.x), e.y & kMetaDefLeft, p.missing) ? 1u : 0u);
          e = __ldg(tree + n);
        }
        leaf_val = __uint_as_float(e.x);
        leaf_id = (int32_t)n - 1;
      } else {
        const uint4* tree = p.sparse + p.tree_offsets[t];
        uint32_t n = 0;
        uint4 e = __ldg(tree);
        while (!(e.y & kMetaLeaf)) {
          const int fid = e.y & 0x3FFFFFFF;
          const float v = SMEM_ROWS ? s_rows[fid * kDirectThreads + threadIdx.x] : __ldg(grow + fid);
          n = e.z + (branch_right(v, __uint_as_float(e.x), e.y & kMetaDefLeft, p.missing) ? 1u : 0u);
          e = __ldg(tree + n);
        }
        leaf_val = __uint_as_float(e.x);
        leaf_id = (int32_t)n;
      }
      acc += leaf_val;
      if (p.leaf_ids) p.leaf_ids[row * (size_t)p.num_trees + t] = leaf_id;
    }
    p.preds[row] = apply_transform(acc, p.tf);
    return;
  }
  // Small batches: kDirectGroup trees are walked together: the level step is branch-free (a walk on its leaf
  // re-reads it), so the dependent L2 round trips of the walks overlap and lanes whose walk ended
  // early do not idle through the longest one -- few warps are resident, so latency is what counts.
  for (int t0 = 0; t0 < p.num_trees; t0 += kDirectGroup) {
    uint32_t n[kDirectGroup];
    const void* tree[kDirectGroup];
    uint32_t ex[kDirectGroup], ey[kDirectGroup], ez[kDirectGroup];
#pragma unroll
    for (int g = 0; g < kDirectGroup; ++g) {
      const int t = min(t0 + g, p.num_trees - 1);          // past the end: repeat the last tree, drop its value
      if (!SPARSE) {
        const uint2* tr = p.heap + (int64_t)t * p.tree_stride;
        tree[g] = tr;
        n[g] = 1;
        const uint2 e = __ldg(tr + 1);
        ex[g] = e.x; ey[g] = e.y; ez[g] = 0;
      } else {
        const uint4* tr = p.sparse + p.tree_offsets[t];
        tree[g] = tr;
        n[g] = 0;
        const uint4 e = __ldg(tr);
        ex[g] = e.x; ey[g] = e.y; ez[g] = e.z;
      }
    }
    for (;;) {
      bool live = false;
#pragma unroll
      for (int g = 0; g < kDirectGroup; ++g) {
        const bool leaf = (ey[g] & kMetaLeaf) != 0;
        const int fid = leaf ? 0 : (int)(ey[g] & 0x3FFFFFFF);
        const float v = SMEM_ROWS ? s_rows[fid * kDirectThreads + threadIdx.x] : __ldg(grow + fid);
        const bool c = branch_right(v, __uint_as_float(ex[g]), ey[g] & kMetaDefLeft, p.missing);
        if (!SPARSE) {
          n[g] = leaf ? n[g] : 2 * n[g] + (c ? 1u : 0u);
          const uint2 e = __ldg(static_cast<const uint2*>(tree[g]) + n[g]);
          ex[g] = e.x; ey[g] = e.y;
        } else {
          n[g] = leaf ? n[g] : ez[g] + (c ? 1u : 0u);
          const uint4 e = __ldg(static_cast<const uint4*>(tree[g]) + n[g]);
          ex[g] = e.x; ey[g] = e.y; ez[g] = e.z;
        }
        live |= !leaf;
      }
      if (!__any_sync(__activemask(), live)) break;
    }
#pragma unroll
    for (int g = 0; g < kDirectGroup; ++g) {
      if (t0 + g >= p.num_trees) break;
      acc += __uint_as_float(ex[g]);                       // tree order: the CPU's float sum, bit for bit
      if (p.leaf_ids) p.leaf_ids[row * (size_t)p.num_trees + t0 + g] = SPARSE ? (int32_t)n[g] : (int32_t)n[g] - 1;
    }
  }
  p.preds[row] = apply_transform(acc, p.tf);
}

// Sparse forests, full machine, 8-byte nodes: the lanes of a warp walk the SAME tree (their loads share
// the top levels' sectors), each in its own while loop, trees in order.  Tried and dropped
// (profiles/r01_sparse_notes.md): free-running lanes (every lane in its own tree: full lane
// utilisation, but 32 sectors per load instead of ~5 -- 1.36x slower).
template <bool SMEM_ROWS>
__global__ void __launch_bounds__(kDirectThreads) k_sparse8(DirectParams p) {
  extern __shared__ float s_rows[];  // SMEM_ROWS: [cols][kDirectThreads]
  const size_t block_row0 = (size_t)blockIdx.x * kDirectThreads;
  const size_t row = block_row0 + threadIdx.x;
  if (SMEM_ROWS) {
    const size_t nrows = min((size_t)kDirectThreads, p.rows - block_row0);
    const size_t total = nrows * (size_t)p.cols;
    const float* src = p.data + block_row0 * (size_t)p.cols;
    for (size_t w = threadIdx.x; w < total; w += kDirectThreads) {
      const int r = (int)(w / p.cols), c = (int)(w % p.cols);
      s_rows[c * kDirectThreads + r] = __ldg(src + w);
    }
    __syncthreads();
  }
  if (row >= p.rows) return;
  const float* grow = p.data + row * (size_t)p.cols;
  const uint32_t fid_mask = (1u << p.fid_bits) - 1u;
  float acc = 0.0f;
  for (int t = 0; t < p.num_trees; ++t) {
    const uint2* tree = p.sparse8 + p.tree_offsets[t];
    uint32_t n = 0;
    uint2 e = __ldg(tree);
    while (!(e.y & kMetaLeaf)) {
      const uint32_t fid = e.y & fid_mask;
      const float v = SMEM_ROWS ? s_rows[fid * kDirectThreads + threadIdx.x] : __ldg(grow + fid);
      n += ((e.y & 0x3FFFFFFFu) >> p.fid_bits) +
           (branch_right(v, __uint_as_float(e.x), e.y & kMetaDefLeft, p.missing) ? 1u : 0u);
      e = __ldg(tree + n);
    }
    acc += __uint_as_float(e.x);
    if (p.leaf_ids) p.leaf_ids[row * (size_t)p.num_trees + t] = (int32_t)n;
  }
  p.preds[row] = apply_transform(acc, p.tf);
}

int launch_direct(tahoe_forest* f, const PredictArgs& a, bool rows_in_smem, int* launches) {
  if (a.rows == 0) return TAHOE_OK;
  DirectParams p{};
  p.data = a.data;
  p.rows = a.rows;
  p.cols = f->params.num_cols;
  p.num_trees = f->params.num_trees;
  p.missing = f->params.missing;
  p.heap = f->d_heap;
  p.tree_stride = (int64_t)1 << (f->depth + 1);
  p.sparse = f->d_sparse;
  p.sparse8 = f->d_sparse8;
  p.fid_bits = f->host.sparse_fid_bits;
  p.tree_offsets = f->d_tree_offsets;
  p.preds = a.out;     // one pass: the kernel's only store is the final score
  p.leaf_ids = a.leaf_ids;
  p.tf = Transform{f->params.output, (float)(a.avg_trees ? a.avg_trees : f->params.num_trees), f->params.threshold,
                   f->params.global_bias};
  if (a.tree_hi > a.tree_lo) {   // one class of a class forest (sparse layouts: trees are addressed through tree_offsets)
    if (!f->sparse || a.leaf_ids) { set_error("tree ranges need a sparse forest and no leaf ids"); return TAHOE_ERR_INVALID; }
    p.tree_offsets += a.tree_lo;
    p.num_trees = a.tree_hi - a.tree_lo;
  }
  const size_t blocks = (a.rows + kDirectThreads - 1) / kDirectThreads;
  if (blocks > 0x7FFFFFFFull) { set_error("too many rows for one launch"); return TAHOE_ERR_INVALID; }
  const size_t smem = rows_in_smem ? (size_t)p.cols * kDirectThreads * sizeof(float) : 0;
  if (smem > (size_t)f->smem_optin) {
    set_error("SHARED_DATA_PER_THREAD: a block's rows do not fit shared memory");
    return TAHOE_ERR_NOT_APPLICABLE;
  }
  auto go = [&](auto kern) -> int {
    if (smem > 48 * 1024)
      TAHOE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)blocks, kDirectThreads, smem, a.stream>>>(p);
    TAHOE_CUDA(cudaGetLastError());
    ++*launches;
    return TAHOE_OK;
  };
  // few rows: not enough warps to hide the L2 round trips -> 4 walks per thread
  int group = a.rows <= (size_t)f->sm_count * 1024 ? 4 : 1;
  if (f->tune.direct_group) group = f->tune.direct_group == 4 ? 4 : 1;
  if (f->sparse && group == 1 && f->d_sparse8)   // full machine, 8-byte nodes when the fields fit
    return rows_in_smem ? go(k_sparse8<true>) : go(k_sparse8<false>);
  if (group == 4) {
    if (f->sparse) return rows_in_smem ? go(k_direct<true, true, 4>) : go(k_direct<true, false, 4>);
    return rows_in_smem ? go(k_direct<false, true, 4>) : go(k_direct<false, false, 4>);
  }
  if (f->sparse) return rows_in_smem ? go(k_direct<true, true, 1>) : go(k_direct<true, false, 1>);
  return rows_in_smem ? go(k_direct<false, true, 1>) : go(k_direct<false, false, 1>);
}

}  // namespace tahoe
