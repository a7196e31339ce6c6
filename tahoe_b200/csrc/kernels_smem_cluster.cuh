// kernels_smem_cluster.cuh -- k_smem_forest_cluster: SPLIT_SHARED_FOREST (reference K5, Struct.h:1414-1606) as a
// thread-block-CLUSTER pipeline, for forests whose trees are so deep that a chunk holds one traversal group
// (config 3: 4 trees of 32 KB) and the forest therefore takes hundreds of chunk passes, each of which
// re-streams every rank tile from HBM (ncu, round 2: 291 GB of DRAM reads per 10M-row launch, 72 % of peak).
//
// The S CTAs of a cluster hold S CONSECUTIVE chunks of the forest and work on the SAME tile at about the same time:
//   * the tile leaves HBM once per cluster (the S requests for it meet in L2), so DRAM traffic for tiles drops by S;
//   * every CTA walks its own G trees for the tile; the row sums stay in TREE ORDER, bit for bit: CTA 0 adds its
//     leaf values to the carried partial sum and hands the result to CTA 1 through distributed shared memory
//     (st.async: each store completes 4 bytes of a transaction on CTA 1's mbarrier), CTA 1 adds ITS leaf values
//     (already looked up: the walk does not depend on the sum) and hands on, the last CTA stores the row.
// One pass of the cluster = S chunks; the cluster barrier replaces __syncthreads between passes.
#pragma once
#include "kernels_smem_traverse.cuh"

namespace tahoe {

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ uint32_t cluster_id_x() {
  uint32_t r; asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r)); return r;
}
__device__ __forceinline__ uint32_t nclusters_x() {
  uint32_t r; asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r)); return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// the shared::cluster address of `local` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local, uint32_t rank) {
  uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank)); return r;
}
// One float into another CTA's shared memory, completing 4 bytes of a transaction on an mbarrier of THAT CTA: data and
// signal travel together, so the hand-over needs no fence (a release.cluster arrive compiles to MEMBAR.ALL.GPU and
// stalls the warp behind every outstanding global access: measured 2x-5x slower than no cluster at all).
__device__ __forceinline__ void st_async_f32(uint32_t raddr, float v, uint32_t rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f32 [%0], %1, [%2];" ::"r"(raddr), "f"(v), "r"(rbar)
               : "memory");
}
// plain remote arrive (the consumer-release of a pipeline: the reads it orders have been consumed already)
__device__ __forceinline__ void mbar_arrive_remote(uint32_t raddr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(raddr) : "memory");
}
// Shared-memory map (byte offsets from smem[]), the same in every CTA of the cluster (mapa relies on it):
//   [0, tile_region)              slot-major tile region, two buffers per warp (slot = warp*2 + b), slot_stride bytes
//                                 each with the columns padded to a power of two: the first two PADDING column rows of
//                                 a slot (128 bytes each, at (cols + j)*128) are the warp's ring of FOUR hand-over rows
//   [.., + trees_per_chunk*W*4)   this CTA's chunk
//   [.., bars)                    mbarriers: chunk | full[warp][2] | acc[warp][4] | accfree[warp][4]
//
// Data flows ONE way, CTA r -> CTA r+1, and nothing on the critical path waits for a neighbour:
//   * every CTA brings the tile into its own buffers itself (plain bulk copy, as k_smem_forest does).  The CTAs of
//     a cluster ask for the same tile within a microsecond of each other, so HBM is read once per cluster and L2
//     serves the others (a multicast copy was tried: the leader must then wait for every CTA's buffer release
//     before each refill, two remote hops per tile -- 58.4 ms against 51.2 without clusters);
//   * the 32 sums of tile k go into ring row k & 3 of the next CTA with st.async, which completes a transaction on
//     that CTA's acc barrier (data and signal together: no fence); the receiver looks at them only after its own
//     walk of tile k, by which time they have long arrived;
//   * back-pressure: the receiver arrives on accfree[k & 3] of the sender when it has consumed row k & 3; the sender
//     waits for that before it hands over tile k + 4 -- four tiles of slack, so it does not stall in steady state.
template <int DEPTH, int G>
__global__ void __launch_bounds__(kSmemMaxThreads, 1) k_smem_forest_cluster(const __grid_constant__ SmemParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr uint32_t W = 1u << (DEPTH + 1);      // words per tree
  const TopWords* no_tw = nullptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  const uint32_t rank = cluster_ctarank(), S = cluster_nctarank();
  const uint32_t tile_bytes = (uint32_t)p.cols * 128u;
  const uint32_t forest_off = p.tile_region;
  const uint32_t bars_off = forest_off + (uint32_t)p.trees_per_chunk * W * 4u;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t bar_chunk = s_base + bars_off;
  const uint32_t bar_full = bar_chunk + 8 + (uint32_t)warp * 16;                                   // + b*8
  const uint32_t bar_acc = bar_chunk + 8 + (uint32_t)nwarps * 16 + (uint32_t)warp * 32;            // + q*8
  const uint32_t bar_accfree = bar_chunk + 8 + (uint32_t)nwarps * 48 + (uint32_t)warp * 32;        // + q*8
  // ring row q of this warp: slot (q & 1), padding column row (q >> 1)
  auto ring_row = [&](uint32_t q) {
    return s_base + (uint32_t)(warp * 2 + (q & 1u)) * p.slot_stride + ((uint32_t)p.cols + (q >> 1)) * 128u + (uint32_t)lane * 4u;
  };

  if (threadIdx.x == 0) {
    mbar_init(bar_chunk, 1);
    for (int i = 0; i < nwarps * 2; ++i) mbar_init(bar_chunk + 8 + i * 8, 1);                      // full
    for (int i = 0; i < nwarps * 4; ++i) {
      mbar_init(bar_chunk + 8 + nwarps * 16 + i * 8, 1);                                          // acc: the local expect_tx arrive
      mbar_init(bar_chunk + 8 + nwarps * 48 + i * 8, 1);                                          // accfree: one arrive from the CTA after
    }
    fence_mbar_init();
  }
  cluster_sync_all();            // every CTA's barriers exist before a neighbour stores into or arrives on them

  const size_t unit_global = (size_t)cluster_id_x() * nwarps + warp;      // tiles are dealt to (cluster, warp)
  const size_t unit_stride = (size_t)nclusters_x() * nwarps;
  uint32_t full_phase = 0, acc_phase = 0, accfree_phase = 0, chunk_phase = 0;   // bit i = parity to wait for on barrier i
  const int passes = (p.num_chunks + (int)S - 1) / (int)S;
  uint32_t k_abs = 0;            // tiles this warp has handled since the kernel began: ring position and phases run across passes

  auto load_tile = [&](size_t tile, int b) {
    if (lane == 0) {
      fence_proxy_async();
      mbar_expect_tx(bar_full + b * 8, tile_bytes);
      bulk_g2s(s_base + (uint32_t)(warp * 2 + b) * p.slot_stride, p.qtiles + tile * (size_t)p.cols * 32, tile_bytes, bar_full + b * 8);
    }
  };

  for (int pass = 0; pass < passes; ++pass) {
    const int chunk = pass * (int)S + (int)rank;                 // this CTA's chunk in this pass (may be past the end)
    const bool have = chunk < p.num_chunks;
    const int tree0 = chunk * p.trees_per_chunk;
    // ---- stage this CTA's chunk.  The hand-over of pass n+1 (rank 0 reads preds[]) needs the last CTA's stores of
    // pass n: the cluster barrier orders them (release / acquire at cluster scope covers global memory)
    if (pass) cluster_sync_all();
    if (threadIdx.x == 0 && have) {
      fence_proxy_async();
      const uint32_t bytes = (uint32_t)G * W * 4u;
      mbar_expect_tx(bar_chunk, bytes);
      const uint8_t* src = (const uint8_t*)(p.qwords + (size_t)tree0 * W);
      uint32_t done = 0;
      while (done < bytes) {
        const uint32_t piece = min(bytes - done, 65536u);
        bulk_g2s(s_base + forest_off + done, src + done, piece, bar_chunk);
        done += piece;
      }
    }
    size_t tile = unit_global;
    int b = 0;
    if (tile < p.num_tiles) load_tile(tile, 0);
    if (have) { mbar_wait(bar_chunk, chunk_phase); chunk_phase ^= 1; }

    const bool first = p.first_pass && pass == 0, last = p.last_pass && pass == passes - 1;
    uint32_t flag_next = 0;
    float acc_next = 0.0f;
    if (tile < p.num_tiles) {
      flag_next = __ldg(p.tile_flags + tile);
      const size_t r0 = tile * 32 + lane;
      // (__ldcg: the partial sums were stored by ANOTHER SM, the cluster's last CTA -- read them from L2, never from this SM's L1)
      if (rank == 0 && !first && r0 < p.rows) acc_next = __ldcg(p.preds + r0);
    }
    for (; tile < p.num_tiles; tile += unit_stride, ++k_abs) {
      const size_t next = tile + unit_stride;
      const uint32_t q = k_abs & 3u;
      const bool has_missing = flag_next != 0;
      float acc = acc_next;
      __syncwarp();                                     // all lanes finished reading the buffer about to be refilled
      if (rank > 0 && lane == 0) mbar_expect_tx(bar_acc + q * 8, 128u);   // the 32 sums the CTA before hands over for this tile
      if (next < p.num_tiles) {
        load_tile(next, b ^ 1);
        flag_next = __ldg(p.tile_flags + next);
        const size_t rn = next * 32 + lane;
        acc_next = (rank == 0 && !first && rn < p.rows) ? __ldcg(p.preds + rn) : 0.0f;
      }
      mbar_wait(bar_full + b * 8, (full_phase >> b) & 1u);
      full_phase ^= 1u << b;

      const size_t row = tile * 32 + lane;
      const bool live = row < p.rows;
      const uint32_t slot_lane = (uint32_t)(warp * 2 + b) * p.slot_stride + lane * 4u;
      float leaf[G];
      if (have) {
        uint32_t a[G];
        const uint32_t tree_abs = s_base + forest_off;
        if (has_missing) traverse_group<DEPTH, G, true, 0>(tree_abs, s_base, slot_lane, p.col_mask, p.miss_code, a, *no_tw, 0);
        else traverse_group<DEPTH, G, false, 0>(tree_abs, s_base, slot_lane, p.col_mask, p.miss_code, a, *no_tw, 0);
#pragma unroll
        for (int g = 0; g < G; ++g) leaf[g] = __uint_as_float(lds_u32(a[g]));
      }
      // ---- the partial sum arrives from the CTA before (CTA 0 carries it from preds[]), in tree order
      if (rank > 0) {
        mbar_wait(bar_acc + q * 8, (acc_phase >> q) & 1u);
        acc_phase ^= 1u << q;
        acc = __uint_as_float(lds_u32(ring_row(q)));
        __syncwarp();                                   // every lane has its sum: ring row q may be written again
        if (lane == 0) mbar_arrive_remote(map_to_cta(bar_accfree + q * 8, rank - 1));
      }
      if (have) {
#pragma unroll
        for (int g = 0; g < G; ++g) acc += leaf[g];     // a padding tree adds +0.0f: the sum keeps its bits
      }
      if (rank + 1 < S) {
        // ring row q of the next CTA was last used four tiles ago: wait until it has been consumed
        if (k_abs >= 4) {
          mbar_wait(bar_accfree + q * 8, (accfree_phase >> q) & 1u);
          accfree_phase ^= 1u << q;
        }
        st_async_f32(map_to_cta(ring_row(q), rank + 1), acc, map_to_cta(bar_acc + q * 8, rank + 1));
      } else if (live) {
        if (last) p.out[row] = apply_transform(acc, p.tf);
        else p.preds[row] = acc;
      }
      b ^= 1;
    }
  }
  cluster_sync_all();            // no CTA exits while a neighbour may still arrive on its barriers or store into it
}

}  // namespace tahoe
