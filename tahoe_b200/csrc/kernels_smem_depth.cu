// kernels_smem_depth.cu -- the instantiations of k_smem_forest for ONE tree depth.
// Compiled 14 times (-DTAHOE_DEPTH=1..14, tahoe_b200/build.py) so the builds run in parallel; each
// object exports launch_smem_depth_<N>(), which picks the instantiation the handle's plan names:
// trees per group (1..6), TMA double-buffered or register-staged tiles.
#include "kernels_smem_traverse.cuh"
#include "kernels_smem_deep.cuh"
#include "kernels_smem_cluster.cuh"

// deep walks in flight per lane (k_smem_forest_deep; each holds a 32-byte block in registers), for the
// 256-thread and the 512-thread instantiations.  Config 4 on B200, 1M rows, top depth 6 / 7:
// 16 warps in pairs: 1 -> 9.60 / 9.79 ms, 2 -> 9.00 / 9.38, 3 -> 9.18 / 9.95;
// 8 warps: 2 -> 11.21 / 10.54, 3 -> 11.12 / 10.87, 4 -> 11.76 / 11.55
#ifndef TAHOE_DEEP_WALKS
#define TAHOE_DEEP_WALKS 2
#endif
#ifndef TAHOE_DEEP_WALKS_512
#define TAHOE_DEEP_WALKS_512 2
#endif

#ifndef TAHOE_DEPTH
#error "compile with -DTAHOE_DEPTH=<1..14>"
#endif
#define TAHOE_CAT2(a, b) a##b
#define TAHOE_CAT(a, b) TAHOE_CAT2(a, b)

namespace tahoe {

int TAHOE_CAT(launch_smem_depth_, TAHOE_DEPTH)(tahoe_forest* f, const SmemParams& p, const TopWords& tw,
                                              unsigned grid, cudaStream_t stream) {
  constexpr int D = TAHOE_DEPTH;
  auto go = [&](auto kern) -> int {
    TAHOE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f->smem_bytes));
    kern<<<grid, (unsigned)f->warps * 32u, f->smem_bytes, stream>>>(p, tw);
    TAHOE_CUDA(cudaGetLastError());
    return TAHOE_OK;
  };
#if TAHOE_DEPTH <= 12
  // sparse forests (QT layout): dense tops of depth D, deep walks pipelined through L2 (kernels_smem_deep.cuh)
  if (f->sparse) {
    constexpr int K = TAHOE_DEEP_WALKS, K5 = TAHOE_DEEP_WALKS_512;
    constexpr int P8 = kStagedMaxCols / 4, P16 = kStagedMaxColsDeep / 4;
    // rows of <= 32 columns: one warp per tile, 8 staging registers; wider rows: 8 warps with 16
    // staging registers, or 16 warps in pairs that share a tile (8 each)
    const bool wide = p.cols > kStagedMaxCols;
    const bool pair = wide && f->warps > 8;
    const int v = !wide ? 0 : (pair ? 2 : 1);
    if (p.leaf_ids)
      return v == 0 ? go(k_smem_forest_deep<D, 2, true, P8, K5, false>)
           : v == 1 ? go(k_smem_forest_deep<D, 2, true, P16, K, false>)
                    : go(k_smem_forest_deep<D, 2, true, P8, K5, true>);
#define TAHOE_DEEP_CASE(G)                                                   \
  case G:                                                                    \
    return v == 0 ? go(k_smem_forest_deep<D, G, false, P8, K5, false>)       \
         : v == 1 ? go(k_smem_forest_deep<D, G, false, P16, K, false>)       \
                  : go(k_smem_forest_deep<D, G, false, P8, K5, true>);
    switch (f->group) {
      TAHOE_DEEP_CASE(2) TAHOE_DEEP_CASE(4) TAHOE_DEEP_CASE(6)
      default: break;
    }
#undef TAHOE_DEEP_CASE
    set_error("unsupported traversal group size for a sparse forest");
    return TAHOE_ERR_INVALID;
  }
#else
  if (f->sparse) { set_error("sparse tops deeper than 12 levels are not built"); return TAHOE_ERR_INVALID; }
#endif
#if TAHOE_DEPTH >= 10
  // cluster pipeline (deep trees: one traversal group per chunk): S CTAs hold S consecutive chunks, tiles by multicast
  if (p.cluster > 1 && !p.leaf_ids) {
    auto go_cluster = [&](auto kern) -> int {
      TAHOE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f->smem_bytes));
      cudaLaunchConfig_t cfg{};
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = (unsigned)p.cluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
      cfg.blockDim = dim3((unsigned)f->warps * 32u); cfg.dynamicSmemBytes = f->smem_bytes; cfg.stream = stream;
      cfg.attrs = attr; cfg.numAttrs = 1;
      cfg.gridDim = dim3((unsigned)p.cluster);
      int max_clusters = 0;
      if (cudaOccupancyMaxActiveClusters(&max_clusters, kern, &cfg) != cudaSuccess) { cudaGetLastError(); return TAHOE_ERR_NOT_APPLICABLE; }
      if (max_clusters < 1) return TAHOE_ERR_NOT_APPLICABLE;   // (the caller falls back to the plain kernel)
      const size_t want = (p.num_tiles + f->warps - 1) / f->warps;
      cfg.gridDim = dim3((unsigned)(std::min<size_t>((size_t)max_clusters, want) * (size_t)p.cluster));
      TAHOE_CUDA(cudaLaunchKernelEx(&cfg, kern, p));
      return TAHOE_OK;
    };
    int rc = TAHOE_ERR_NOT_APPLICABLE;
    switch (f->group) {
      case 2: rc = go_cluster(k_smem_forest_cluster<D, 2>); break;
      case 3: rc = go_cluster(k_smem_forest_cluster<D, 3>); break;
      case 4: rc = go_cluster(k_smem_forest_cluster<D, 4>); break;
      case 5: rc = go_cluster(k_smem_forest_cluster<D, 5>); break;
      case 6: rc = go_cluster(k_smem_forest_cluster<D, 6>); break;
      default: break;
    }
    // a device (or partition) that cannot co-schedule the cluster: the plain kernel below does the same work
    if (rc != TAHOE_ERR_NOT_APPLICABLE) return rc;
  }
#endif
  // the leaf-id instantiation is the checking path: one tree at a time keeps the build small
  if (p.leaf_ids)
    return f->staged ? go(k_smem_forest<D, 1, true, true, 0>)
                     : (f->single ? go(k_smem_forest<D, 1, true, false, 0, true>) : go(k_smem_forest<D, 1, true, false, 0>));
  // p.top_levels: 2 = levels 0-1 of every tree from the kernel-parameter table, 0 = every level from shared memory
#define TAHOE_G_CASE(G)                                                                                         \
  case G:                                                                                                       \
    if (f->single)                                                                                              \
      return p.top_levels >= 2 ? go(k_smem_forest<D, G, false, false, 2, true>)                                 \
                               : go(k_smem_forest<D, G, false, false, 0, true>);                                \
    if (p.top_levels >= 2)                                                                                      \
      return f->staged ? go(k_smem_forest<D, G, false, true, 2>) : go(k_smem_forest<D, G, false, false, 2>);    \
    return f->staged ? go(k_smem_forest<D, G, false, true, 0>) : go(k_smem_forest<D, G, false, false, 0>);
  switch (f->group) {
    TAHOE_G_CASE(1) TAHOE_G_CASE(2) TAHOE_G_CASE(3) TAHOE_G_CASE(4) TAHOE_G_CASE(5) TAHOE_G_CASE(6)
    default: break;
  }
#undef TAHOE_G_CASE
  set_error("unsupported traversal group size");
  return TAHOE_ERR_INVALID;
}

}  // namespace tahoe
