// packer.cu -- forest loader / packer (host side).
//
// B200-first redesign of dense_adaptive_forest::init + swap_child
// (Struct.h:1712-1986) and encode_node_adaptive (Struct.h:77-98):
//   * probability reorder: a node whose right child is heavier gets its two
//     subtrees swapped and an exchange bit (same rule, Struct.h:1781-1786),
//   * a REAL simhash over (level, feature) tokens orders similar trees next to
//     each other (the reference hashes uninitialised buffers, Struct.h:1854-1872),
//   * node-major interleaved layout [node][tree] for the lane=tree kernel
//     (Struct.h:1911-1923),
//   * the reference's adaptive 1/2/4-byte feature-id width (Struct.h:1827-1852)
//     becomes a rank-quantised 4-byte node: the float threshold is replaced by
//     its rank among the forest's distinct thresholds on that column, so a whole
//     node (rank, column, default direction) is one shared-memory word and the
//     branch test is one unsigned compare.
#include <algorithm>
#include <limits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>

#include "forest.h"

namespace tahoe {

static inline uint32_t f2u(float x) { uint32_t u; std::memcpy(&u, &x, 4); return u; }

static inline uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

static int ceil_log2(uint64_t v) { int b = 0; while ((1ull << b) < v) ++b; return b; }

// Reference width rule, Struct.h:1831-1852: fid_len = (log2(max_fid)+3)/8.
static int width_rule(int max_fid) {
  if (max_fid <= 0) return 0;  // log(0): undefined in the reference
  double len = (std::log((double)max_fid) / std::log(2.0) + 3.0) / 8.0;
  if (len > 0.0 && len < 1.0) return 1;
  if (len >= 1.0 && len < 2.0) return 2;
  if (len >= 2.0 && len < 4.0) return 4;
  return 0;
}

// ---- dense ---------------------------------------------------------------------
struct TreeScratch {
  std::vector<float> val, weight;
  std::vector<uint32_t> meta;
  std::vector<int32_t> orig;
};

static void swap_subtrees(TreeScratch& t, int64_t a, int64_t b, int64_t npt) {
  // iterative level-by-level swap of the two complete subtrees rooted at a and b
  int64_t width = 1;
  while (a < npt) {
    for (int64_t i = 0; i < width; ++i) {
      std::swap(t.val[a + i], t.val[b + i]);
      std::swap(t.weight[a + i], t.weight[b + i]);
      std::swap(t.meta[a + i], t.meta[b + i]);
      std::swap(t.orig[a + i], t.orig[b + i]);
    }
    a = 2 * a + 1;
    b = 2 * b + 1;
    width *= 2;
  }
}

static int build_quant(tahoe_forest* f, const tahoe_dense_node_t* nodes);
static void plan_sparse_smem(tahoe_forest* f, const tahoe_sparse_node_t* nodes);
static void plan_sparse_qtop(tahoe_forest* f);
static double plan_chunks_for(tahoe_forest* f, int depth, bool deep);

int pack_dense(tahoe_forest* f, const tahoe_dense_node_t* nodes) {
  const int T = f->params.num_trees, d = f->params.depth, C = f->params.num_cols;
  if (T < 0 || d < 0 || d > 26 || C < 1) { set_error("invalid forest parameters"); return TAHOE_ERR_INVALID; }
  const int64_t npt = ((int64_t)1 << (d + 1)) - 1;
  const int64_t W = (int64_t)1 << (d + 1);
  HostPack& h = f->host;
  f->depth = d;

  // ---- reachability, validation, statistics
  std::vector<uint8_t> reach((size_t)npt);
  int max_fid = 0;
  int64_t real_nodes = 0;
  double depth_acc = 0.0;
  for (int t = 0; t < T; ++t) {
    const tahoe_dense_node_t* tr = nodes + (int64_t)t * npt;
    std::fill(reach.begin(), reach.end(), 0);
    reach[0] = 1;
    for (int64_t k = 0; k < npt; ++k) {
      if (!reach[k]) continue;
      ++real_nodes;
      const bool leaf = tr[k].bits < 0;
      if (leaf) continue;
      const int fid = tr[k].bits & 0x3FFFFFFF;
      if (2 * k + 2 >= npt) { set_error("reachable internal node on the bottom level"); return TAHOE_ERR_INVALID; }
      if (fid >= C) { set_error("feature id >= num_cols"); return TAHOE_ERR_INVALID; }
      if (fid >= (1 << 29)) { set_error("feature id needs more than 29 bits"); return TAHOE_ERR_INVALID; }
      max_fid = std::max(max_fid, fid);
      reach[2 * k + 1] = reach[2 * k + 2] = 1;
      int level = 0; for (int64_t q = k + 1; q > 1; q >>= 1) ++level;
      depth_acc += std::ldexp(1.0, -level);
    }
  }
  f->num_nodes = real_nodes;
  f->node_width = width_rule(max_fid);
  f->mean_depth = T ? depth_acc / T : 0.0;

  // ---- H8: 1-based heap, 8 B/node
  h.heap.assign((size_t)T * W, make_uint2(0, kMetaLeaf));
  for (int t = 0; t < T; ++t) {
    const tahoe_dense_node_t* tr = nodes + (int64_t)t * npt;
    uint2* out = h.heap.data() + (int64_t)t * W;
    for (int64_t k = 0; k < npt; ++k) out[k + 1] = make_uint2(f2u(tr[k].val), (uint32_t)tr[k].bits);
  }

  // ---- I8: probability reorder + simhash order + node-major interleave
  h.tpad = ((T + 31) / 32) * 32;
  if (h.tpad == 0) h.tpad = 32;
  const bool build_inter = (double)npt * h.tpad * 12.0 <= 6.0e9;
  h.simhash.assign(T, 0);
  std::vector<TreeScratch> trees(build_inter ? T : 0);
  int64_t exchanged = 0;
  for (int t = 0; t < T; ++t) {
    const tahoe_dense_node_t* tr = nodes + (int64_t)t * npt;
    // simhash over (level, fid) tokens of reachable internal nodes
    int cnt[64] = {0};
    std::fill(reach.begin(), reach.end(), 0);
    reach[0] = 1;
    for (int64_t k = 0; k < npt; ++k) {
      if (!reach[k] || tr[k].bits < 0) continue;
      reach[2 * k + 1] = reach[2 * k + 2] = 1;
      int level = 0; for (int64_t q = k + 1; q > 1; q >>= 1) ++level;
      uint64_t hsh = mix64(((uint64_t)level << 32) ^ (uint64_t)(tr[k].bits & 0x3FFFFFFF));
      // shallow nodes decide more traffic: weight tokens by 2^-(level/2)
      int wgt = std::max(1, 16 >> (level / 2));
      for (int b = 0; b < 64; ++b) cnt[b] += ((hsh >> b) & 1) ? wgt : -wgt;
    }
    uint64_t sh = 0;
    for (int b = 0; b < 64; ++b) if (cnt[b] > 0) sh |= 1ull << b;
    h.simhash[t] = sh;
    if (!build_inter) continue;
    TreeScratch& s = trees[t];
    s.val.resize(npt); s.weight.resize(npt); s.meta.resize(npt); s.orig.resize(npt);
    for (int64_t k = 0; k < npt; ++k) {
      s.val[k] = tr[k].val;
      s.weight[k] = tr[k].weight;
      uint32_t b = (uint32_t)tr[k].bits;
      s.meta[k] = (b & kMetaFidMask) | (b & kMetaDefLeft) | (b & kMetaLeaf);
      s.orig[k] = (int32_t)k;
    }
    // bottom-up child swap, Struct.h:1775-1825
    for (int lv = f->tune.no_reorder ? -1 : d - 1; lv >= 0; --lv) {
      for (int64_t k = ((int64_t)1 << lv) - 1; k < ((int64_t)1 << (lv + 1)) - 1; ++k) {
        if (s.meta[k] & kMetaLeaf) continue;
        if (s.weight[2 * k + 1] < s.weight[2 * k + 2]) {
          s.meta[k] |= kMetaExchange;
          swap_subtrees(s, 2 * k + 1, 2 * k + 2, npt);
          ++exchanged;
        }
      }
    }
  }
  f->exchanged = exchanged;
  h.tree_order.assign(h.tpad, -1);
  {
    std::vector<int> order(T);
    std::iota(order.begin(), order.end(), 0);
    if (!f->tune.no_simhash)
      std::stable_sort(order.begin(), order.end(),
                       [&](int a, int b) { return h.simhash[a] < h.simhash[b]; });
    for (int i = 0; i < T; ++i) h.tree_order[i] = order[i];
  }
  if (build_inter) {
    h.inter.assign((size_t)npt * h.tpad, make_uint2(0, kMetaLeaf));  // padding slot = leaf 0.0
    h.orig_ids.assign((size_t)npt * h.tpad, 0);
    for (int slot = 0; slot < T; ++slot) {
      const TreeScratch& s = trees[h.tree_order[slot]];
      for (int64_t k = 0; k < npt; ++k) {
        h.inter[(size_t)k * h.tpad + slot] = make_uint2(f2u(s.val[k]), s.meta[k]);
        h.orig_ids[(size_t)k * h.tpad + slot] = s.orig[k];
      }
    }
  }
  if (build_inter) plan_shared_data(f);
  return build_quant(f, nodes);
}

// ---- SHARED_DATA plan + B32 blocks ----------------------------------------------------------
// A team of up to 4 warps shares a row (two row buffers per team, lane = tree slot).  Shared memory
// is split between row buffers (teams in flight) and the top levels of the forest; below them a walk
// pays one L2 round trip per level (~800 cycles of a dependent chain, a 32-byte sector per 8-byte
// node once the lanes' paths have parted).  B32 halves both: the levels below the top are stored as
// two-level blocks {root, left child, right child} in one sector, so one trip decides two levels.
void plan_shared_data(tahoe_forest* f) {
  HostPack& h = f->host;
  const int d = f->depth, tpad = h.tpad;
  const size_t row_bytes = (size_t)f->params.num_cols * sizeof(float);
  const size_t row_stride = (row_bytes + 15) & ~(size_t)15;
  const int slot_groups = tpad / 32;
  f->sd_team_warps = slot_groups >= 4 ? 4 : (slot_groups >= 2 ? 2 : 1);
  // Wide rows (config 1: 12 KB): the row buffers bound the teams in flight, and the kernel streams rows from
  // HBM.  Teams of two warps with ONE buffer each put 15 rows in flight per SM instead of 8; a team's own
  // load latency is exposed but the other teams cover it (B200, 50K x 3072: 0.168 -> 0.142 ms; one-warp
  // teams 0.152, two-warp teams with two buffers 0.165 -- gpurun_out/r2d_c1.log, profiles/r02_c1_plans.md).
  const bool wide_rows = row_stride >= 4096 && slot_groups >= 2;
  if (wide_rows) f->sd_team_warps = 2;
  if (const char* e = std::getenv("TAHOE_SD_TEAM")) f->sd_team_warps = std::max(1, std::min(f->sd_team_warps, std::atoi(e)));
  const int per_warp = (slot_groups + f->sd_team_warps - 1) / f->sd_team_warps;   // slot groups each warp walks
  f->sd_group = per_warp >= 4 ? 4 : (per_warp >= 2 ? 2 : 1);
  f->sd_row_bufs = wide_rows ? 1 : 2;
  if (const char* e = std::getenv("TAHOE_SD_BUFS")) f->sd_row_bufs = std::atoi(e) == 1 ? 1 : 2;
  const size_t per_team = (size_t)f->sd_row_bufs * row_stride + 64 + 16;    // row buffers, partials, mbarriers
  f->sd_teams = 0;
  f->sd_blocked = false;
  if (per_team + 128 > (size_t)f->smem_optin) return;  // SHARED_DATA not applicable: two rows do not fit
  const int max_teams = std::min(15, 32 / f->sd_team_warps);          // named barriers 1..15; 1024 threads
  // the shared-memory copy of the top levels is [node][stride] uint2: the blocked walk skips the padding slots
  // (stride = trees: config 1 keeps 7 levels of its 100 trees in 102 KB instead of 130), the plain walk keeps
  // stride = tpad (bank = slot: conflict-free)
  const bool tight = !std::getenv("TAHOE_SD_LOOSE_TOP");
  auto top_stride = [&](bool blocked) { return (size_t)(blocked && tight ? f->params.num_trees : tpad); };
  auto top_bytes = [&](int top, bool blocked) {
    return top ? ((((size_t)1 << top) - 1) * top_stride(blocked) * sizeof(uint2) + 15) / 16 * 16 : (size_t)0;
  };
  auto teams_for = [&](int top, bool blocked) {
    return (int)std::min<size_t>((size_t)max_teams, ((size_t)f->smem_optin - 256 - top_bytes(top, blocked)) / per_team);   // (128 of slack + the round-up of the team region)
  };
  // blocked form: top levels K with (d + 1 - K) even and K <= d - 1, blocks small enough to keep
  const char* eb = std::getenv("TAHOE_SD_BLOCKED");
  const bool want_blocked = d >= 2 && !(eb && std::atoi(eb) == 0);
  int best_top = 0;
  bool best_blocked = false;
  double best_rate = 0.0;
  // Below the top a blocked walk makes (d - top) decisions, two per L2 trip, the last trip's children being the
  // bottom-level leaves: (d + 1 - top) must be even.  (A 64-byte tail block -- two decisions and the leaf in one
  // trip, for the even splits -- was built in round 2 and is slower at every split of config 1:
  // profiles/r02_c1_walk_cycles.md.)
  auto trips_for = [&](int top) { return (d + 1 - top) / 2; };
  auto blocked_bytes = [&](int top) {
    size_t roots = 0;
    for (int L = top; L < d; L += 2) roots += (size_t)1 << L;
    return roots * 32 * tpad;
  };
  auto blocked_ok = [&](int top) {
    return want_blocked && top >= 1 && top <= d - 1 && !((d + 1 - top) & 1) && row_bytes <= kSdOffMask &&
           blocked_bytes(top) <= ((size_t)512 << 20);
  };
  for (int blocked = 0; blocked <= (want_blocked ? 1 : 0); ++blocked) {
    for (int top = blocked ? 1 : 0; top <= (blocked ? d - 1 : d + 1); ++top) {
      if (top_bytes(top, blocked != 0) + per_team + 128 > (size_t)f->smem_optin) break;
      if (blocked && !blocked_ok(top)) continue;
      // What the plans of config 1 measure (profiles/r02_c1_walk_cycles.md): the SM's rows per cycle hardly depend on
      // the teams once ~10 of them keep it busy (a walk takes longer in proportion), an L2 trip costs a little more
      // than two shared-memory levels, and fewer trips win: 7 levels + 1 trip with 10 teams beats 5 + 2 with 15.
      const double rate = std::min(teams_for(top, blocked != 0), 10) /
                          (60.0 * top + 160.0 * (blocked ? trips_for(top) : d + 1 - top) + 100.0);
      if (rate > best_rate) { best_rate = rate; best_top = top; best_blocked = blocked != 0; }
    }
  }
  if (const char* e = std::getenv("TAHOE_SD_TOP_LEVELS")) {
    int top = std::max(0, std::min(std::atoi(e), d + 1));
    if (top_bytes(top, blocked_ok(top)) + per_team + 128 <= (size_t)f->smem_optin) {
      best_top = top;
      best_blocked = blocked_ok(top);
    }
  }
  f->sd_top = best_top;
  f->sd_blocked = best_blocked;
  f->sd_teams = teams_for(best_top, best_blocked);
  f->sd_top_stride = (int)top_stride(best_blocked);
  f->sd_smem = ((size_t)f->sd_teams * per_team + 127) / 128 * 128 + top_bytes(best_top, best_blocked) + 128;
  if (!best_blocked) return;

  const int K = best_top;
  const int64_t npt = ((int64_t)1 << (d + 1)) - 1;
  size_t roots = 0;
  for (int L = K; L < d; L += 2) roots += (size_t)1 << L;
  h.sd_top_nodes.assign((((size_t)1 << K) - 1) * tpad, make_uint2(0, kMetaLeaf));
  h.sd_blocks.assign(roots * tpad * 2, make_uint4(0, 0, 0, 0));
  h.sd_orig_bottom.assign(((size_t)1 << d) * tpad, 0);
  std::vector<uint2> pd((size_t)npt);
  std::vector<int32_t> po((size_t)npt);
  for (int slot = 0; slot < tpad; ++slot) {
    for (int64_t k = 0; k < npt; ++k) {       // leaves pushed down: everything under a leaf is that leaf
      uint2 node = h.inter[(size_t)k * tpad + slot];
      int32_t o = h.orig_ids[(size_t)k * tpad + slot];
      if (k > 0 && (pd[(k - 1) / 2].y & kMetaLeaf)) { node = pd[(k - 1) / 2]; o = po[(k - 1) / 2]; }
      pd[k] = node;
      po[k] = o;
    }
    // the walk's own node form (kSd* in forest.h): byte offset of the feature, the exchange bit folded into the
    // outcome of a missing value and into a flip of the comparison -- a decision is six instructions shorter
    for (int64_t k = 0; k < npt; ++k) {
      const uint32_t y = pd[k].y;
      if (y & kMetaLeaf) { pd[k].y = kSdLeaf; continue; }          // pushed down: both children are this leaf
      const bool exch = (y & kMetaExchange) != 0, def_left = (y & kMetaDefLeft) != 0;
      pd[k].y = ((!def_left) != exch ? kSdMissRight : 0u) | (exch ? kSdFlip : 0u) | ((y & kMetaFidMask) * 4u);
    }
    for (int64_t k = 0; k < ((int64_t)1 << K) - 1; ++k) h.sd_top_nodes[(size_t)k * tpad + slot] = pd[k];
    size_t base = 0;
    for (int L = K; L < d; L += 2) {
      const int64_t first = ((int64_t)1 << L) - 1;
      for (int64_t q = 0; q < ((int64_t)1 << L); ++q) {
        const int64_t r = first + q;
        uint4* blk = h.sd_blocks.data() + ((base + q) * tpad + slot) * 2;
        blk[0] = make_uint4(pd[r].x, pd[r].y, pd[2 * r + 1].x, pd[2 * r + 1].y);
        blk[1] = make_uint4(pd[2 * r + 2].x, pd[2 * r + 2].y, 0u, 0u);
      }
      base += (size_t)1 << L;
    }
    const int64_t bottom = ((int64_t)1 << d) - 1;
    for (int64_t q = 0; q < ((int64_t)1 << d); ++q) h.sd_orig_bottom[(size_t)q * tpad + slot] = po[bottom + q];
  }
  const size_t top_nodes = ((size_t)1 << K) - 1, stride = (size_t)f->sd_top_stride;
  h.sd_top_image.assign((top_nodes * stride + 1) / 2 * 2, make_uint2(0, kSdLeaf));
  for (size_t k = 0; k < top_nodes; ++k)
    for (size_t slot = 0; slot < stride; ++slot) h.sd_top_image[k * stride + slot] = h.sd_top_nodes[k * tpad + slot];
}

// ---- Q4: rank-quantised heap ------------------------------------------------------
// Rank tables shared by the dense Q4 layout and the sparse quantised tops (QT): per column the sorted
// distinct thresholds (k_quantize turns a feature into rank(v) = #{thr <= v}) and the bucket index
// that narrows the search.  Sets rank_bits / col_bits / rank_shift; false when rank + column + the
// 7 lane/word bits do not fit a 32-bit node word.
static bool build_rank_tables(tahoe_forest* f, std::vector<std::vector<float>>& thr) {
  HostPack& h = f->host;
  const int C = f->params.num_cols;
  h.qoffsets.assign(C + 1, 0);
  size_t max_n = 0;
  for (int c = 0; c < C; ++c) {
    std::sort(thr[c].begin(), thr[c].end());
    thr[c].erase(std::unique(thr[c].begin(), thr[c].end()), thr[c].end());  // == merges -0/+0
    max_n = std::max(max_n, thr[c].size());
    h.qoffsets[c + 1] = h.qoffsets[c] + (int32_t)thr[c].size();
  }
  const int rb = std::max(1, ceil_log2(max_n + 2));  // ranks 0..n, sentinel code 2^rb-1 > n
  const int cb = std::max(1, ceil_log2((uint64_t)C));
  if (rb + cb + kQuantColShift > 32) return false;
  h.qtable.clear();
  h.qtable.reserve(h.qoffsets[C]);
  for (int c = 0; c < C; ++c) h.qtable.insert(h.qtable.end(), thr[c].begin(), thr[c].end());
  // bucket LUT (k_quantize): bucket(x) = clamp(int((x - min) * scale), 0, B-1) is monotone in x, so
  // thresholds in lower buckets are < v and thresholds in higher buckets are > v; only v's own
  // bucket needs a search.  Host and device evaluate bucket() with the same IEEE float ops.
  {
    // about two buckets per threshold of the busiest column; k_quantize keeps ONE column's index in
    // shared memory at a time, so the bound is per column
    int B = 16;
    while (B < kQuantMaxBuckets && (size_t)B < 2 * max_n) B <<= 1;
    h.qbuckets = B;
    h.qlut_min.assign(C, 0.0f);
    h.qlut_scale.assign(C, 0.0f);
    h.qlut.assign((size_t)C * (B + 1), 0);
    for (int c = 0; c < C; ++c) {
      const std::vector<float>& tv = thr[c];
      int32_t* lut = h.qlut.data() + (size_t)c * (B + 1);
      if (tv.empty()) continue;
      const float mn = tv.front(), range = tv.back() - tv.front();
      float scale = 0.0f;
      if (range > 0.0f && std::isfinite(range)) {
        scale = (float)B / range;
        if (!std::isfinite(scale)) scale = 0.0f;
      }
      h.qlut_min[c] = mn;
      h.qlut_scale[c] = scale;
      for (float t : tv) {
        const float x = (t - mn) * scale;
        int b = x >= (float)B ? B - 1 : (int)x;   // tv is sorted, so x >= 0
        if (b < 0) b = 0;
        if (b > B - 1) b = B - 1;
        lut[b + 1] += 1;
      }
      for (int b = 0; b < B; ++b) lut[b + 1] += lut[b];
    }
  }
  f->rank_bits = rb;
  f->col_bits = cb;
  f->rank_shift = 32 - rb;
  return true;
}

static int build_quant(tahoe_forest* f, const tahoe_dense_node_t* nodes) {
  const int T = f->params.num_trees, d = f->depth, C = f->params.num_cols;
  HostPack& h = f->host;
  f->quant = false;
  if (d < 1 || d > kQuantMaxDepth || C > kQuantMaxCols || T < 1) return TAHOE_OK;
  const int64_t npt = ((int64_t)1 << (d + 1)) - 1;
  const int64_t W = (int64_t)1 << (d + 1);
  const int64_t L = (int64_t)1 << d;

  // state per 1-based heap position: 0 = reachable internal, >0 = under/at leaf (origin+1)
  std::vector<int32_t> origin((size_t)W);
  std::vector<std::vector<float>> thr(C);
  for (int t = 0; t < T; ++t) {
    const tahoe_dense_node_t* tr = nodes + (int64_t)t * npt;
    origin[1] = (tr[0].bits < 0) ? 1 : 0;
    for (int64_t i = 1; i < L; ++i) {
      int32_t o = origin[i];
      if (o == 0) {
        float v = tr[i - 1].val;
        if (!std::isnan(v)) thr[tr[i - 1].bits & 0x3FFFFFFF].push_back(v);
        for (int c = 0; c < 2; ++c) origin[2 * i + c] = (tr[2 * i + c - 1].bits < 0) ? (int32_t)(2 * i + c) : 0;
      } else {
        origin[2 * i] = origin[2 * i + 1] = o;
      }
    }
  }
  if (!build_rank_tables(f, thr)) return TAHOE_OK;  // rank + column code does not fit one word: no Q4 layout
  f->quant = true;            // plan_chunks needs it; cleared again if no tree fits shared memory
  f->qdepth = d;
  plan_chunks(f);
  if (!f->quant) return TAHOE_OK;   // not even one tree fits beside the sample tiles
  // col_shift was chosen with the plan: the column field of a node word is always the byte offset of
  // the column row in the plan's tile region
  f->col_mask = ((1u << f->col_bits) - 1u) << f->col_shift;
  f->miss_code = ((1u << f->rank_bits) - 1u) << f->rank_shift;

  h.qwords.assign((size_t)T * W, 0u);
  h.leaf_origin.assign((size_t)T * L, 0);
  for (int t = 0; t < T; ++t) {
    const tahoe_dense_node_t* tr = nodes + (int64_t)t * npt;
    uint32_t* out = h.qwords.data() + (int64_t)t * W;
    int32_t* lo = h.leaf_origin.data() + (int64_t)t * L;
    origin[1] = (tr[0].bits < 0) ? 1 : 0;
    for (int64_t i = 1; i < W; ++i) {
      int32_t o = origin[i];
      if (i < L) {
        if (o == 0) {
          const int fid = tr[i - 1].bits & 0x3FFFFFFF;
          const float v = tr[i - 1].val;
          uint32_t k;
          if (std::isnan(v)) k = (uint32_t)thr[fid].size();  // val >= NaN is never true
          else k = (uint32_t)(std::lower_bound(thr[fid].begin(), thr[fid].end(), v) - thr[fid].begin());
          out[i] = (k << f->rank_shift) | ((uint32_t)fid << f->col_shift) |
                   ((tr[i - 1].bits & 0x40000000) ? 1u : 0u);
          for (int c = 0; c < 2; ++c) origin[2 * i + c] = (tr[2 * i + c - 1].bits < 0) ? (int32_t)(2 * i + c) : 0;
        } else {
          out[i] = 0u;  // below an early leaf: both children lead to the same value
          origin[2 * i] = origin[2 * i + 1] = o;
        }
      } else {
        // bottom slot: o > 0 always (validated: reachable bottom nodes are leaves)
        out[i] = f2u(tr[o - 1].val);
        lo[i - L] = o - 1;
      }
    }
  }
  return TAHOE_OK;
}

// ---- sparse (compact) forests --------------------------------------------------------
// Replaces sparse_forest::init (Struct.h:2327-2345).  Nodes stay breadth-first per
// tree with the two children adjacent, so both children of a node share one 32-byte
// sector; the 12-byte reference node is padded to 16 B for one aligned 128-bit load.
int pack_sparse(tahoe_forest* f, const tahoe_sparse_node_t* nodes, const float* weights,
                const int64_t* tree_offsets) {
  (void)weights;  // children are adjacent: the probability swap has nothing to win here
  const int T = f->params.num_trees, C = f->params.num_cols;
  if (T < 0 || C < 1 || !tree_offsets) { set_error("invalid sparse forest parameters"); return TAHOE_ERR_INVALID; }
  HostPack& h = f->host;
  h.tree_offsets.assign(tree_offsets, tree_offsets + T + 1);
  const int64_t total = tree_offsets[T];
  h.sparse.resize((size_t)total);
  int max_fid = 0, max_level = 0;
  double depth_acc = 0.0;
  std::vector<int32_t> level;
  for (int t = 0; t < T; ++t) {
    const int64_t base = tree_offsets[t], n = tree_offsets[t + 1] - base;
    if (n < 1) { set_error("empty tree"); return TAHOE_ERR_INVALID; }
    level.assign((size_t)n, -1);
    level[0] = 0;
    for (int64_t k = 0; k < n; ++k) {
      const tahoe_sparse_node_t& nd = nodes[base + k];
      h.sparse[base + k] = make_uint4(f2u(nd.val), (uint32_t)nd.bits, (uint32_t)nd.left_idx, 0u);
      if (nd.bits < 0 || level[k] < 0) continue;  // leaf, or unreachable
      const int fid = nd.bits & 0x3FFFFFFF;
      if (fid >= C) { set_error("feature id >= num_cols"); return TAHOE_ERR_INVALID; }
      if (nd.left_idx <= k || (int64_t)nd.left_idx + 1 >= n) {
        set_error("sparse child index out of range (children must follow their parent)");
        return TAHOE_ERR_INVALID;
      }
      max_fid = std::max(max_fid, fid);
      level[nd.left_idx] = level[nd.left_idx + 1] = level[k] + 1;
      max_level = std::max(max_level, level[k] + 1);
      depth_acc += std::ldexp(1.0, -level[k]);
    }
  }
  // S8: children follow their parent in breadth-first order, so the child index is stored as a
  // delta from the node's own index; with the feature id below it the node is 8 bytes, a child pair
  // 16, and one 32-byte sector holds two pairs
  {
    const int fb = std::max(1, ceil_log2((uint64_t)std::max(max_fid + 1, 2)));
    const uint64_t max_delta = (1ull << (30 - fb)) - 1;
    bool fits = fb <= 20;
    for (int t = 0; t < T && fits; ++t) {
      const int64_t base = tree_offsets[t], n = tree_offsets[t + 1] - base;
      for (int64_t k = 0; k < n; ++k) {
        const tahoe_sparse_node_t& nd = nodes[base + k];
        if (nd.bits < 0) continue;
        if (nd.left_idx <= k || (uint64_t)(nd.left_idx - k) > max_delta) { fits = false; break; }
      }
    }
    if (fits) {
      h.sparse8.resize((size_t)total);
      h.sparse_fid_bits = fb;
      for (int t = 0; t < T; ++t) {
        const int64_t base = tree_offsets[t], n = tree_offsets[t + 1] - base;
        for (int64_t k = 0; k < n; ++k) {
          const tahoe_sparse_node_t& nd = nodes[base + k];
          uint32_t y;
          if (nd.bits < 0) y = kMetaLeaf;
          else y = ((uint32_t)nd.bits & kMetaDefLeft) | ((uint32_t)(nd.left_idx - k) << fb) |
                   ((uint32_t)nd.bits & ((1u << fb) - 1u));
          h.sparse8[base + k] = make_uint2(f2u(nd.val), y);
        }
      }
    }
  }
  f->sparse = true;
  f->depth = max_level;
  f->num_nodes = total;
  f->node_width = width_rule(max_fid);
  f->mean_depth = T ? depth_acc / T : 0.0;
  f->quant = false;
  plan_sparse_smem(f, nodes);
  plan_sparse_qtop(f);
  return TAHOE_OK;
}

// ---- T8: sparse forests through shared memory ------------------------------------------------
// Deep unbalanced trees do not fit shared memory, but their walks are short: most end within a few
// levels of the root.  Each tree is renumbered breadth-first (the top of the tree comes first) and
// the first `cut` nodes of every tree of a chunk are kept in shared memory; a walk that steps past
// them is finished in L2 on the full breadth-first copy.  The cut balances chunk passes (each one
// re-streams the sample tiles) against those deep walks.
static void plan_sparse_smem(tahoe_forest* f, const tahoe_sparse_node_t* nodes) {
  HostPack& h = f->host;
  const int T = f->params.num_trees, C = f->params.num_cols;
  f->sp_cut = f->sp_tpc = f->sp_chunks = 0;
  f->sp_smem = 0;
  if (h.sparse8.empty() || T < 1 || C > 64) return;   // the 8-byte node fields must fit (top form: 6-bit fid)
  const int fb = h.sparse_fid_bits;
  const uint64_t max_delta = (1ull << (30 - fb)) - 1;
  const size_t tile_bytes = (size_t)kSparseWarps * C * 128;      // one float tile per warp
  const size_t fixed = tile_bytes + (size_t)kSparseWarps * kSparsePend * 8 + 1024;
  if (fixed + 4096 > (size_t)f->smem_optin) return;
  const size_t tree_budget = (size_t)f->smem_optin - fixed;

  // breadth-first renumbering; level and reach probability (children equally likely) per node
  const int64_t total = h.tree_offsets[T];
  h.sp_full.assign((size_t)total, make_uint2(0, kMetaLeaf));
  h.sp_orig.assign((size_t)total, 0);
  std::vector<int32_t> newidx, order;
  int64_t max_n = 0;
  // over_by_cut[j]: expected deep walks per (row, forest) for cut = 32 << j
  const int kCuts = 7;   // cut <= 2048: the shared-memory node form keeps delta * 8 in 15 bits
  double over_by_cut[kCuts] = {0};
  for (int t = 0; t < T; ++t) {
    const int64_t base = h.tree_offsets[t], n = h.tree_offsets[t + 1] - base;
    max_n = std::max(max_n, n);
    newidx.assign((size_t)n, -1);
    order.clear();
    order.push_back(0);
    newidx[0] = 0;
    for (size_t q = 0; q < order.size(); ++q) {
      const int32_t k = order[q];
      const tahoe_sparse_node_t& nd = nodes[base + k];
      if (nd.bits < 0) continue;
      if (order.size() + 2 > (size_t)n) { h.sp_full.clear(); h.sp_orig.clear(); return; }   // shared children: not a tree
      for (int c = 0; c < 2; ++c) {
        newidx[nd.left_idx + c] = (int32_t)order.size();
        order.push_back(nd.left_idx + c);
      }
    }
    // reach probabilities in breadth-first numbering
    std::vector<double> pr(order.size(), 0.0);
    pr[0] = 1.0;
    for (size_t q = 0; q < order.size(); ++q) {
      const tahoe_sparse_node_t& nd = nodes[base + order[q]];
      uint32_t y;
      if (nd.bits < 0) {
        y = kMetaLeaf;
      } else {
        const int64_t nl = newidx[nd.left_idx];
        if ((uint64_t)(nl - (int64_t)q) > max_delta) { h.sp_full.clear(); h.sp_orig.clear(); return; }
        y = ((uint32_t)nd.bits & kMetaDefLeft) | ((uint32_t)(nl - (int64_t)q) << fb) |
            ((uint32_t)nd.bits & ((1u << fb) - 1u));
        pr[nl] = pr[nl + 1] = 0.5 * pr[q];
        for (int j = 0; j < kCuts; ++j) {
          const int64_t cut = (int64_t)32 << j;
          if ((int64_t)q < cut && nl + 1 >= cut) over_by_cut[j] += (nl >= cut ? 2 : 1) * 0.5 * pr[q];
        }
      }
      h.sp_full[base + q] = make_uint2(f2u(nd.val), y);
      h.sp_orig[base + q] = order[q];
    }
  }
  // cut: a power of two (or, capped, the longest tree), cheapest under
  //   chunks * (one pass of the tile stream + per-tile fixed cost) + deep walks * their L2 cost
  int best_j = -1;
  double best_cost = 0.0;
  int want = 0;
  if (const char* e = std::getenv("TAHOE_SPARSE_CUT")) want = std::atoi(e);
  for (int j = 0; j < kCuts; ++j) {
    const int64_t cut = (int64_t)32 << j;
    if (want && cut != want) continue;
    const size_t tpc = tree_budget / ((size_t)cut * 8);
    if (tpc < 1) break;
    const double chunks = std::ceil((double)T / (double)std::min<size_t>(tpc, (size_t)T));
    const double cost = chunks * 1.0e-10 + over_by_cut[j] * 3.1e-11;   // B200 fit: profiles/r01_sparse_notes.md
    if (best_j < 0 || cost < best_cost) { best_j = j; best_cost = cost; }
    if (cut >= max_n) break;                       // every tree already whole
  }
  if (best_j < 0) { h.sp_full.clear(); h.sp_orig.clear(); return; }
  const int64_t cut = (int64_t)32 << best_j;
  f->sp_cut = (int)cut;
  f->sp_tpc = (int)std::min<size_t>(tree_budget / ((size_t)cut * 8), (size_t)T);
  if (const char* e = std::getenv("TAHOE_SPARSE_TPC")) f->sp_tpc = std::max(1, std::min(f->sp_tpc, std::atoi(e)));
  f->sp_chunks = (T + f->sp_tpc - 1) / f->sp_tpc;
  f->sp_overflow = over_by_cut[best_j] / T;
  f->sp_smem = fixed + (size_t)f->sp_tpc * cut * 8;
  h.sp_top.assign((size_t)f->sp_chunks * f->sp_tpc * cut, make_uint2(0, 1u));   // padding: leaf 0.0
  for (int t = 0; t < T; ++t) {
    const int64_t base = h.tree_offsets[t], n = h.tree_offsets[t + 1] - base;
    // shared-memory form of a top node (kernels_sparse_smem.cu):
    //   y = (delta * 8) << 17 | fid << 11 | def_left << 1 | leaf   (delta <= cut + 1 < 4096)
    for (int64_t q = 0; q < std::min<int64_t>(n, cut); ++q) {
      const uint2 e = h.sp_full[base + q];
      uint32_t y = 1u;
      if (!(e.y & kMetaLeaf)) {
        const uint32_t delta = (e.y & 0x3FFFFFFFu) >> fb, fid = e.y & ((1u << fb) - 1u);
        y = ((delta * 8u) << 17) | (fid << 11) | ((e.y & kMetaDefLeft) ? 2u : 0u);
      }
      h.sp_top[(size_t)t * cut + q] = make_uint2(e.x, y);
    }
  }
}

// ---- QT: sparse forests as rank-quantised dense tops + deep walks in L2 ----------------------
// The top `D` levels of every sparse tree are laid out as a dense Q4 heap (leaves pushed down, one
// 32-bit word per node: the six-instruction visit of k_smem_forest applies unchanged).  A bottom slot
// holds either the float bits of the leaf the walk has reached or, when the walk is still on an
// internal node at level D, a continuation marker: a negative quiet NaN whose low 22 bits are the
// node's breadth-first index relative to the first node of its chunk's first tree.  The kernel parks
// such walks (lane-private rings) and every lane advances its own, one 32-byte two-level block per traversal
// group, in L2 on `sp_qblocks` -- nodes in rank form {rank word, block delta | leaf}, compared against the
// same rank tile in shared memory.  Thresholds of ALL internal nodes go into the
// rank tables, so one quantise pass serves both parts.
constexpr uint32_t kContMask = 0xFFC00000u;
// one parked deep walk, in shared-memory node-visit units: its L2 round trips hide behind the traversal
// groups (k_smem_forest_deep), what is left is the instructions and the 32-byte gathers of ~2.3 two-level
// blocks.  Config 4 on B200, 1M rows, 16 warps in pairs: top depth 5 / 6 / 7 -> 10.0 / 9.2 / 9.95 ms with
// 540 / 385 / 273 deep walks per row
constexpr double kDeepWalkVisits = 15.0;
// one traversal group of the deep kernel beyond its visits (deep step, parking), in node-visit units:
// 6 trees per group 9.2 ms, 4 per group 10.7 ms on the same forest
constexpr double kDeepGroupVisits = 18.0;

static void plan_sparse_qtop(tahoe_forest* f) {
  HostPack& h = f->host;
  const int T = f->params.num_trees, C = f->params.num_cols;
  f->qt = false;
  f->sp_use_qt = false;
  f->qt_depth = 0;
  if (h.sp_full.empty() || T < 1 || C > kQuantMaxCols || f->depth < 1) return;
  if (const char* e = std::getenv("TAHOE_SPARSE_QTOP")) if (std::atoi(e) == 0) return;
  const int fb = h.sparse_fid_bits;
  const uint32_t fmask = (1u << fb) - 1u;
  const int64_t total = h.tree_offsets[T];
  auto u2f = [](uint32_t u) { float x; std::memcpy(&x, &u, 4); return x; };

  std::vector<std::vector<float>> thr(C);
  int64_t max_n = 0;
  for (int t = 0; t < T; ++t) max_n = std::max(max_n, h.tree_offsets[t + 1] - h.tree_offsets[t]);
  if (max_n >= (int64_t)1 << 22) return;                       // the marker keeps 22 bits of node index
  for (int64_t q = 0; q < total; ++q) {
    const uint2 e = h.sp_full[q];
    if (e.y & kMetaLeaf) {
      if ((e.x & kContMask) == kContMask) return;              // a leaf value that looks like a marker
    } else {
      const float v = u2f(e.x);
      if (!std::isnan(v)) thr[e.y & fmask].push_back(v);
    }
  }
  auto give_up = [&]() {
    h.qtable.clear(); h.qoffsets.clear(); h.qlut.clear(); h.qlut_min.clear(); h.qlut_scale.clear();
    h.qwords.clear(); h.leaf_origin.clear(); h.sp_qblocks.clear(); h.sp_qblock_orig.clear(); h.sp_qblock_offsets.clear();
    f->qt = false; f->sp_use_qt = false; f->qt_depth = 0; f->trees_per_chunk = 0; f->num_chunks = 0;
  };
  if (!build_rank_tables(f, thr)) { give_up(); return; }

  // walks still on an internal node at level D, per (row, forest), children taken as equally likely
  const int max_d = std::min(std::min(f->depth, kQuantMaxDepth), 12);
  std::vector<double> over((size_t)max_d + 2, 0.0);
  {
    std::vector<double> pr;
    std::vector<int32_t> lev;
    for (int t = 0; t < T; ++t) {
      const int64_t base = h.tree_offsets[t], n = h.tree_offsets[t + 1] - base;
      pr.assign((size_t)n, 0.0);
      lev.assign((size_t)n, 0);
      pr[0] = 1.0;
      for (int64_t q = 0; q < n; ++q) {
        const uint2 e = h.sp_full[base + q];
        if ((e.y & kMetaLeaf) || pr[q] == 0.0) continue;
        if (lev[q] <= max_d) over[lev[q]] += pr[q];
        const int64_t nl = q + ((e.y & 0x3FFFFFFFu) >> fb);
        pr[nl] = pr[nl + 1] = 0.5 * pr[q];
        lev[nl] = lev[nl + 1] = lev[q] + 1;
      }
    }
  }
  int want_d = 0;
  if (const char* e = std::getenv("TAHOE_QTOP_DEPTH")) want_d = std::atoi(e);
  int best_d = 0;
  double best_cost = 0.0;
  for (int D = 1; D <= max_d; ++D) {
    if (want_d && D != want_d) continue;
    const double c = plan_chunks_for(f, D, true);
    if (c < 0.0) continue;
    const double cost = c + over[D] * kDeepWalkVisits;
    if (!best_d || cost < best_cost) { best_d = D; best_cost = cost; }
  }
  if (!best_d || plan_chunks_for(f, best_d, true) < 0.0) { give_up(); return; }
  const int D = best_d;
  if (const char* e = std::getenv("TAHOE_QTOP_TPC")) {          // test knob: force several chunks
    const int G = std::max(f->group, 1);
    const int cap = std::max(G, std::atoi(e) / G * G);
    if (cap < f->trees_per_chunk) {
      f->trees_per_chunk = cap;
      f->num_chunks = (f->qwords_trees + cap - 1) / cap;
    }
  }
  f->qt = true;
  {
    const char* path = std::getenv("TAHOE_SPARSE_PATH");
    f->sp_use_qt = !(path && std::strcmp(path, "t8") == 0) && !std::getenv("TAHOE_SPARSE_CUT");
  }
  f->qt_depth = D;
  f->qdepth = D;
  f->qt_overflow = over[D] / T;
  f->col_mask = ((1u << f->col_bits) - 1u) << f->col_shift;
  f->miss_code = ((1u << f->rank_bits) - 1u) << f->rank_shift;

  auto rank_word = [&](uint2 e) -> uint32_t {
    const uint32_t fid = e.y & fmask;
    const float v = u2f(e.x);
    const std::vector<float>& tv = thr[fid];
    const uint32_t k = std::isnan(v) ? (uint32_t)tv.size()       // val >= NaN is never true
                                     : (uint32_t)(std::lower_bound(tv.begin(), tv.end(), v) - tv.begin());
    return (k << f->rank_shift) | (fid << f->col_shift) | ((e.y & kMetaDefLeft) ? 1u : 0u);
  };
  const int64_t W = (int64_t)1 << (D + 1), L = (int64_t)1 << D;
  h.qwords.assign((size_t)T * W, 0u);
  h.leaf_origin.assign((size_t)T * L, -1);
  // Below the top: TWO-LEVEL BLOCKS of 32 bytes {root, left child, right child, -}, each node
  // {rank word | leaf value, leaf << 31 | block delta}: one L2 sector decides two levels of a deep
  // walk.  Block roots are the nodes at levels D, D+2, ...; a child's delta is the distance from its
  // block to the block of ITS left child (the right child's block follows it).  Blocks of one tree are
  // numbered breadth-first from the internal nodes at level D, in bottom-slot order.
  h.sp_qblocks.clear();
  h.sp_qblock_orig.clear();
  h.sp_qblock_offsets.assign((size_t)T + 1, 0);
  auto deep_node = [&](int64_t base, int32_t q, uint32_t delta) -> uint2 {
    const uint2 e = h.sp_full[base + q];
    return (e.y & kMetaLeaf) ? make_uint2(e.x, kMetaLeaf) : make_uint2(rank_word(e), delta);
  };
  auto left_child = [&](int64_t base, int32_t q) -> int32_t {
    return q + (int32_t)((h.sp_full[base + q].y & 0x3FFFFFFFu) >> fb);
  };
  std::vector<int32_t> at((size_t)W), roots;
  int64_t chunk_first_block = 0;
  for (int t = 0; t < T; ++t) {
    const int64_t base = h.tree_offsets[t];
    const int64_t block0 = (int64_t)h.sp_qblocks.size() / 2;       // first block of this tree
    h.sp_qblock_offsets[t] = block0;
    if (t % f->trees_per_chunk == 0) chunk_first_block = block0;
    roots.clear();
    uint32_t* out = h.qwords.data() + (int64_t)t * W;
    int32_t* lo = h.leaf_origin.data() + (int64_t)t * L;
    at[1] = 0;
    for (int64_t i = 1; i < W; ++i) {
      const int32_t q = at[i];
      const uint2 e = h.sp_full[base + q];
      const bool leaf = (e.y & kMetaLeaf) != 0;
      if (i < L) {
        if (leaf) {
          out[i] = 0u;                                   // below a leaf: both children lead to it
          at[2 * i] = at[2 * i + 1] = q;
        } else {
          out[i] = rank_word(e);
          const int32_t nl = q + (int32_t)((e.y & 0x3FFFFFFFu) >> fb);
          at[2 * i] = nl;
          at[2 * i + 1] = nl + 1;
        }
      } else if (leaf) {
        out[i] = e.x;
        lo[i - L] = h.sp_orig[base + q];
      } else {
        // the walk goes on below the top, in the block rooted here (index relative to the chunk's first block)
        const int64_t rel = block0 - chunk_first_block + (int64_t)roots.size();
        if (rel >= (int64_t)1 << 22) { give_up(); return; }    // the marker keeps 22 bits
        out[i] = kContMask | (uint32_t)rel;
        roots.push_back(q);
      }
    }
    // breadth-first over block roots: roots[b] is the root of block block0 + b
    for (size_t b = 0; b < roots.size(); ++b) {
      const int32_t q = roots[b];
      const bool leaf = (h.sp_full[base + q].y & kMetaLeaf) != 0;
      uint4 w0 = make_uint4(0u, kMetaLeaf, 0u, kMetaLeaf), w1 = make_uint4(0u, kMetaLeaf, 0u, 0u);
      int32_t orig[3] = {h.sp_orig[base + q], -1, -1};
      if (leaf) {
        w0.x = h.sp_full[base + q].x;
      } else {
        const uint2 r = deep_node(base, q, 0u);
        w0.x = r.x; w0.y = 0u;
        const int32_t kid = left_child(base, q);
        uint2 k[2];
        for (int c = 0; c < 2; ++c) {
          const int32_t qc = kid + c;
          orig[1 + c] = h.sp_orig[base + qc];
          uint32_t delta = 0;
          if (!(h.sp_full[base + qc].y & kMetaLeaf)) {         // its two children root the next blocks
            delta = (uint32_t)(roots.size() - b);
            const int32_t g = left_child(base, qc);
            roots.push_back(g);
            roots.push_back(g + 1);
          }
          k[c] = deep_node(base, qc, delta);
        }
        w0.z = k[0].x; w0.w = k[0].y; w1.x = k[1].x; w1.y = k[1].y;
      }
      h.sp_qblocks.push_back(w0);
      h.sp_qblocks.push_back(w1);
      h.sp_qblock_orig.insert(h.sp_qblock_orig.end(), orig, orig + 3);
    }
    if ((int64_t)h.sp_qblocks.size() / 2 - chunk_first_block >= (int64_t)1 << 30) { give_up(); return; }
  }
  h.sp_qblock_offsets[T] = (int64_t)h.sp_qblocks.size() / 2;
  // the dummy block idle walk slots sit on: a leaf root worth +0.0f
  h.sp_qblocks.push_back(make_uint4(0u, kMetaLeaf, 0u, kMetaLeaf));
  h.sp_qblocks.push_back(make_uint4(0u, kMetaLeaf, 0u, 0u));
  h.sp_qblock_orig.insert(h.sp_qblock_orig.end(), {-1, -1, -1});
}

// ---- shared-memory chunk plan ------------------------------------------------------------
// Shared memory of one persistent CTA (227 KB opt-in, vs the reference's 0.8 * 48 KB,
// Struct.h:215-220): [warp tiles][forest chunk][mbarriers].
// Relative cost of one node visit for a plan, from B200 measurements on the SUSY- and Higgs-shaped
// forests (profiles/r01_smem_plans.log): 16 warps x 6 trees is the unit.  What matters is the
// number of dependent chains in flight per SM (warps x trees per thread); TMA-fed tiles with few
// warps lose the overlap that hides the bulk copies.
double smem_visit_cost(int warps, int G, bool staged) {
  // B200 measurements, round 2 (gpurun_out r2d_plans: config 2 on 5M rows, config 3 on 2M rows, one-copy
  // TMA tiles, parameter-table levels at >= 16 warps), unit = 16 warps x 5-6 trees staged (2.9e-13 s):
  //   chains (warps x trees per thread) in flight per SM   24: 1.75   32: 1.45   40: 1.12   >= 48: 1.0
  //   TMA, 3 trees per group: 1.10 (16 x 3 on config 3); staged G = 4: 1.04 (config 2)
  //   staged 12 warps: 1.03 (config 2: 7.93 vs 7.72 ms), 8 warps: 1.04
  const int chains = warps * G;
  const double by_chains = chains >= 48 ? 1.0 : chains >= 40 ? 1.12 : chains >= 32 ? 1.45 : chains >= 24 ? 1.75
                         : chains >= 16 ? 2.2 : chains >= 12 ? 2.6 : 3.0;
  if (staged) {
    const double by_group = G >= 5 ? 1.0 : G == 4 ? 1.04 : G == 3 ? 1.1 : G == 2 ? 1.2 : 1.4;
    return by_chains * by_group * (warps >= 16 ? 1.0 : (warps >= 12 ? 1.03 : (warps >= 8 ? 1.04 : 1.15)));
  }
  static const double by_group[kMaxGroup + 1] = {0.0, 1.50, 1.25, 1.10, 1.0, 1.0, 1.0};
  return by_chains * by_group[G];
}

// Deep trees cost more per visit: the bottom levels of a 2^12-leaf tree are random words (bank conflicts: 2.9
// shared-memory wavefronts per warp-visit on config 3 against 2.6 on config 2) and the pipe runs at 89 % instead of
// 96 %.  Measured 1.3 at depth 12 relative to depth 10; applied to the visits, not to the tile passes.
double smem_depth_factor(int depth) { return 1.0 + 0.15 * std::max(0, depth - 10); }

// One pass of a row's tile through a chunk, in node-visit units.  `mode` = how the warp's tile buffer is filled:
// 1 staged: the tile moves through the LSU twice (C/4 128-bit loads + stores per lane) plus a fixed part --
//   fitted 14.4 visits at 18 columns, 19.5 at 28.
// 0 two buffers by TMA: the fixed part only when the tile is ONE bulk copy (slot-major region); the column-major
//   (wide) region takes one 128-byte copy per column (fitted 22 visits at 18 columns).
// 2 SINGLE, one buffer by TMA refilled from L2 after the walk: the copy's latency is exposed to the warp and covered
//   by the SM's other walks -- fitted on config 3 (profiles/r02_single_tiles.log): 0.5 visits with 80 walks per SM,
//   7 with 60-64, 14.6 with 48.
double smem_tile_cost(int cols, int mode, int warps, int group, bool wide) {
  if (mode == 1) return 0.256 * (2.0 * cols + 20.0);
  if (mode == 2) return std::max(3.0, 1700.0 / std::max(warps * group, 1) - 20.0);
  return 8.0 + (wide ? 0.8 * cols : 0.0);
}

void plan_chunks(tahoe_forest* f) { plan_chunks_for(f, f->depth, false); }

// The plan for trees of `depth` levels.  deep: the sparse quantised-tops variant -- every warp also
// keeps a list of parked deep walks in shared memory, and register-staged tiles may be up to
// kStagedMaxColsDeep columns wide at <= 8 warps (256 threads: 255 registers each).
// Returns the modelled cost per sample in node-visit units (< 0: nothing fits).
static double plan_chunks_for(tahoe_forest* f, int depth, bool deep) {
  f->trees_per_chunk = 0;
  f->num_chunks = 0;
  f->smem_bytes = 0;
  f->group = 1;
  f->slots = 0;
  f->quant_stride = 1;
  f->quant_smem = 0;
  f->qwords_trees = f->params.num_trees;
  if (!f->quant && !deep) return -1.0;
  const int T = f->params.num_trees, C = f->params.num_cols;
  const size_t tile_bytes = (size_t)C * 128;
  const size_t tree_bytes = ((size_t)1 << (depth + 1)) * 4;
  int want_warps = 0, want_staged = -1, want_group = 0;   // TAHOE_SMEM_PLAN="warps,staged,group"
  if (const char* e = std::getenv("TAHOE_SMEM_PLAN")) std::sscanf(e, "%d,%d,%d", &want_warps, &want_staged, &want_group);

  // every (warps, tile buffering, group) that fits; cost per sample (in node-visit units) = node
  // visits at the plan's rate + one tile pass per chunk
  double best = 0.0;
  bool found = false;
  static const int kWarpChoices[] = {16, 12, 8, 4};   // 12: slot-major tile regions only (the wide field needs 2^k slots)
  for (int warps : kWarpChoices) {
    if (want_warps && warps != want_warps) continue;
    // deep: + lane-private rings of parked walks + the pairs' exchange floats
    const size_t bars = 1024 + (deep ? (size_t)warps * kDeepPend * 128 + 512 : 0);
    // tiles: 0 = two buffers per warp by TMA, 1 = one buffer + the next tile in registers, 2 = one buffer by TMA,
    // refilled after the walk from L2 (SINGLE: slot-major region so that a tile is one bulk copy)
    for (int staged = 0; staged <= (deep ? 1 : 2); ++staged) {
      if (want_staged >= 0 && staged != want_staged) continue;
      if (deep && !staged) continue;                   // k_smem_forest_deep: the next tile always waits in registers
      if (staged == 1 && C > (deep ? kStagedMaxColsDeep : kStagedMaxCols)) continue;
      // deep: current + previous tile per warp; rows wider than 32 columns at 16 warps: per PAIR of warps
      const bool paired = deep && C > kStagedMaxCols && warps > 8;
      const int slots = deep ? (paired ? warps : warps * 2) : warps * (staged ? 1 : 2);
      // wide column field (column-major tile region) when rank + column bits leave room for it,
      // else the compact field with a slot-major region whose columns are padded to a power of two
      const bool pow2_slots = (slots & (slots - 1)) == 0;
      if (deep && !pow2_slots) continue;
      const bool wide = staged != 2 && pow2_slots && f->rank_bits + f->col_bits + kQuantColShift + ilog2_c(slots) <= 32;
      const size_t cols_pad = (size_t)1 << ceil_log2((uint64_t)C);
      const size_t tiles = wide ? (size_t)slots * tile_bytes : (size_t)slots * cols_pad * 128;
      if ((size_t)f->smem_optin < tiles + bars + tree_bytes) continue;
      const size_t avail = (size_t)f->smem_optin - tiles - bars;
      const int max_tpc = (int)std::min<size_t>(avail / tree_bytes, (size_t)T + kMaxGroup);   // padding trees count
      for (int G = 1; G <= std::min(kMaxGroup, max_tpc); ++G) {
        if (want_group && G != want_group) continue;
        if (deep && (G & 1)) continue;                 // the deep variant is built for 2, 4, 6 trees per group
        const int tpad = (T + G - 1) / G * G;          // all-zero trees pad the forest to whole groups
        int tpc = std::min(max_tpc / G * G, tpad);
        const int chunks = (tpad + tpc - 1) / tpc;     // balance: same chunk count, even whole groups
        tpc = ((tpad + chunks - 1) / chunks + G - 1) / G * G;
        double visits = (double)tpad * std::max(depth, 1) * smem_visit_cost(warps, G, staged == 1) * smem_depth_factor(depth);
        // levels 0-1 from the kernel-parameter table (launch_smem_forest: plans of >= 16 warps)
        if (warps >= 16 && !deep) visits *= 1.0 - 0.25 * std::min(2, depth) / std::max(depth, 1);
        // deep: the per-group work, and 8 warps (64 staging registers) cover less of it than 16 (measured 1.17x)
        if (deep) visits = visits * (warps <= 8 ? 1.12 : 1.0) + (double)tpad / G * kDeepGroupVisits;
        // (deep: the tile of a pass is committed from registers under the deep walks of the tile before)
        const double cost = visits + (double)chunks * (deep ? 10.0 : smem_tile_cost(C, staged, warps, G, wide));
        if (!found || cost < best) {
          found = true; best = cost;
          f->warps = warps; f->staged = staged == 1; f->single = staged == 2; f->slots = slots; f->group = G;
          f->qwords_trees = tpad; f->trees_per_chunk = tpc;
          f->num_chunks = (tpad + tpc - 1) / tpc;
          f->smem_bytes = tiles + bars + (size_t)tpc * tree_bytes;
          f->tile_region = (uint32_t)tiles;
          f->tile_col_stride = wide ? 128u * (uint32_t)slots : 128u;
          f->tile_slot_stride = wide ? 128u : (uint32_t)(cols_pad * 128);
          f->col_shift = wide ? kQuantColShift + ilog2_c(slots) : kQuantColShift;
        }
      }
    }
  }
  if (!found) { f->quant = false; return -1.0; }

  // k_quantize: one column's sorted thresholds (every quant_stride-th of them when the column is
  // too long) + its bucket index live in shared memory
  size_t max_n = 0;
  for (int c = 0; c < C; ++c)
    max_n = std::max<size_t>(max_n, (size_t)(f->host.qoffsets[c + 1] - f->host.qoffsets[c]));
  const size_t lut_bytes = ((size_t)f->host.qbuckets + 1) * 4;
  const size_t budget = (size_t)f->smem_optin - 2048 - lut_bytes;
  int S = 1;
  while ((max_n / S + 1) * 4 > budget) S <<= 1;
  if (S == 2) S = 4;   // the in-block count reads whole float4s
  f->quant_stride = S;
  f->quant_smem = lut_bytes + (max_n / S + 1) * 4 + 16;
  return best;
}

static void read_tuning(Tuning& t) {
  if (const char* e = std::getenv("TAHOE_ROW_BATCH_BYTES")) t.row_batch_bytes = (size_t)std::atoll(e);
  if (const char* e = std::getenv("TAHOE_SPARSE_GRID")) t.sparse_grid = std::atoi(e);
  if (const char* e = std::getenv("TAHOE_SPARSE_GROUP")) t.sparse_group = std::atoi(e);
  if (const char* e = std::getenv("TAHOE_DIRECT_GROUP")) t.direct_group = std::atoi(e);
  if (const char* e = std::getenv("TAHOE_TOP_TREES")) t.top_trees = std::atoi(e);
  if (const char* e = std::getenv("TAHOE_TOP_LEVELS")) t.top_levels = std::atoi(e);
  if (const char* e = std::getenv("TAHOE_QUANT_TILES")) t.quant_tiles = std::atoi(e);
  if (const char* e = std::getenv("TAHOE_CLUSTER")) t.cluster = std::atoi(e);
  if (std::getenv("TAHOE_SD_NO_L2_AHEAD")) t.sd_no_l2_ahead = true;
  if (const char* e = std::getenv("TAHOE_NO_REORDER")) t.no_reorder = std::atoi(e) != 0;
  if (const char* e = std::getenv("TAHOE_NO_SIMHASH")) t.no_simhash = std::atoi(e) != 0;
  if (const char* e = std::getenv("TAHOE_HOST_SLAB")) std::sscanf(e, "%d,%d,%d", &t.slab_first, &t.slab_growth, &t.slab_cap);
}

int query_device(tahoe_forest* f) {
  read_tuning(f->tune);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    // No GPU (authoring container): the forest can be packed and inspected, with the plan a B200
    // would get.  Every predict call on such a handle fails with TAHOE_ERR_CUDA -- no CPU path.
    cudaGetLastError();
    f->device = -1;
    f->sm_count = 148;
    f->smem_optin = 232448;
    return TAHOE_OK;
  }
  cudaDeviceProp prop;
  TAHOE_CUDA(cudaGetDevice(&f->device));
  TAHOE_CUDA(cudaGetDeviceProperties(&prop, f->device));
  f->sm_count = prop.multiProcessorCount;
  f->smem_optin = (int)prop.sharedMemPerBlockOptin;
  f->l2_bytes = prop.l2CacheSize;
  return TAHOE_OK;
}

template <typename T>
static int to_device(T** dst, const std::vector<T>& src) {
  *dst = nullptr;
  if (src.empty()) return TAHOE_OK;
  TAHOE_CUDA(cudaMalloc((void**)dst, src.size() * sizeof(T)));
  TAHOE_CUDA(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
  return TAHOE_OK;
}

int upload(tahoe_forest* f) {
  if (f->device < 0) return TAHOE_OK;   // GPU-less host: host-side pack only
  HostPack& h = f->host;
  int rc;
  if ((rc = to_device(&f->d_heap, h.heap))) return rc;
  if ((rc = to_device(&f->d_sparse, h.sparse))) return rc;
  if ((rc = to_device(&f->d_sparse8, h.sparse8))) return rc;
  if ((rc = to_device(&f->d_sp_full, h.sp_full))) return rc;
  if ((rc = to_device(&f->d_sp_orig, h.sp_orig))) return rc;
  if ((rc = to_device(&f->d_sp_top, h.sp_top))) return rc;
  if ((rc = to_device(&f->d_sp_qblocks, h.sp_qblocks))) return rc;
  if ((rc = to_device(&f->d_sp_qblock_orig, h.sp_qblock_orig))) return rc;
  if ((rc = to_device(&f->d_sp_qblock_offsets, h.sp_qblock_offsets))) return rc;
  if ((rc = to_device(&f->d_tree_offsets, h.tree_offsets))) return rc;
  if ((rc = to_device(&f->d_inter, h.inter))) return rc;
  if ((rc = to_device(&f->d_tree_order, h.tree_order))) return rc;
  if ((rc = to_device(&f->d_orig_ids, h.orig_ids))) return rc;
  if ((rc = to_device(&f->d_sd_top_image, h.sd_top_image))) return rc;
  if ((rc = to_device(&f->d_sd_blocks, h.sd_blocks))) return rc;
  if ((rc = to_device(&f->d_sd_orig_bottom, h.sd_orig_bottom))) return rc;
  if (!h.qwords.empty()) {   // padded with all-zero trees to whole traversal groups
    const size_t W = (size_t)1 << (f->qdepth + 1);
    const size_t words = (size_t)f->qwords_trees * W;
    TAHOE_CUDA(cudaMalloc((void**)&f->d_qwords, words * 4));
    TAHOE_CUDA(cudaMemset(f->d_qwords, 0, words * 4));
    TAHOE_CUDA(cudaMemcpy(f->d_qwords, h.qwords.data(), h.qwords.size() * 4, cudaMemcpyHostToDevice));
    // levels 0 and 1 of every tree for the kernel-parameter table (padding trees: all zero)
    h.top3.assign((size_t)f->qwords_trees * 3, 0u);
    const size_t real = h.qwords.size() / W;
    for (size_t t = 0; t < real; ++t)
      for (int k = 0; k < 3 && (size_t)(k + 1) < W; ++k) h.top3[t * 3 + k] = h.qwords[t * W + 1 + k];
    if (f->qwords_trees <= kTopTrees) std::memcpy(f->top_words_all.w, h.top3.data(), h.top3.size() * sizeof(uint32_t));
  }
  if (!h.qtable.empty()) {
    // device copy of the threshold tables: every column starts on a 16-byte boundary and is padded
    // with NaN (NaN <= v is false) to whole blocks of quant_stride thresholds plus one, so k_quantize
    // can finish a search with unguarded float4 loads.  d_qoffsets = [C padded starts][C counts].
    const int C = f->params.num_cols;
    const size_t S = (size_t)std::max(f->quant_stride, 4);
    std::vector<float> padded;
    std::vector<int32_t> offs(2 * (size_t)C);
    for (int c = 0; c < C; ++c) {
      const size_t n = (size_t)(h.qoffsets[c + 1] - h.qoffsets[c]);
      offs[c] = (int32_t)padded.size();
      offs[(size_t)C + c] = (int32_t)n;
      padded.insert(padded.end(), h.qtable.begin() + h.qoffsets[c], h.qtable.begin() + h.qoffsets[c + 1]);
      padded.resize(offs[c] + (n / S + 1) * S, std::numeric_limits<float>::quiet_NaN());
    }
    if ((rc = to_device(&f->d_qtable, padded))) return rc;
    if ((rc = to_device(&f->d_qoffsets, offs))) return rc;
  }
  if ((rc = to_device(&f->d_leaf_origin, h.leaf_origin))) return rc;
  if ((rc = to_device(&f->d_qlut_min, h.qlut_min))) return rc;
  if ((rc = to_device(&f->d_qlut_scale, h.qlut_scale))) return rc;
  if ((rc = to_device(&f->d_qlut, h.qlut))) return rc;
  return TAHOE_OK;
}

}  // namespace tahoe
