"""Host-side packer checks (no GPU): every packed layout, traversed by a numpy emulation of
its kernel, must land on the oracle's leaf ids; the probability reorder must match what the
REFERENCE's own packer produced (golden pack_* arrays)."""
import numpy as np
import pytest

from tahoe_b200 import api, synth
from tests import emulate
from tests.helpers import load_case

PACK_CASES = ["tiny", "char_w", "short_w", "ties", "one_feature"]


def _forest(c, **kw):
    return api.Forest.from_dense(c["nodes"], c["trees"], c["depth"], c["cols"],
                                 missing=float(c["missing"]), **kw)


def _tree_order_sum(vals):
    acc = np.zeros(vals.shape[0], np.float32)
    for t in range(vals.shape[1]):
        acc = (acc + vals[:, t]).astype(np.float32)
    return acc


@pytest.mark.parametrize("name", PACK_CASES + ["int_w"])
def test_heap_layout(oracle, name):
    c = load_case(name)
    f = _forest(c)
    want, wids = oracle.predict_dense(c["nodes"], c["trees"], c["depth"], c["data"],
                                      float(c["missing"]), want_leaf_ids=True)
    heap = f.export(api.EXPORT_HEAP_NODES, np.uint32)
    vals, ids = emulate.heap_traverse(heap, c["trees"], c["depth"], c["data"], float(c["missing"]))
    assert np.array_equal(ids, wids)
    assert np.array_equal(_tree_order_sum(vals).view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("name", PACK_CASES)
def test_interleaved_layout_with_exchange_bits(oracle, name):
    c = load_case(name)
    f = _forest(c)
    want, wids = oracle.predict_dense(c["nodes"], c["trees"], c["depth"], c["data"],
                                      float(c["missing"]), want_leaf_ids=True)
    inter = f.export(api.EXPORT_INTERLEAVED, np.uint32)
    oid = f.export(api.EXPORT_ORIG_IDS, np.int32)
    order = f.export(api.EXPORT_TREE_ORDER, np.int32)
    tpad = order.shape[0]
    assert tpad % 32 == 0 and sorted(order[order >= 0].tolist()) == list(range(c["trees"]))
    vals, ids = emulate.interleaved_traverse(inter, oid, order, tpad, c["depth"], c["trees"],
                                             c["data"], float(c["missing"]))
    assert np.array_equal(ids, wids)
    np.testing.assert_allclose(vals.sum(1, dtype=np.float64), want.astype(np.float64), rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("name", PACK_CASES)
@pytest.mark.parametrize("top", [None, 1, 2, 3])
def test_blocked_layout_for_shared_data(oracle, name, top, monkeypatch):
    """B32: pushed-down leaves + two-level 32-byte blocks give the oracle's leaves, for the planned split
    of levels and for forced ones (TAHOE_SD_TOP_LEVELS is read when the forest is packed)."""
    c = load_case(name)
    if top is not None:
        monkeypatch.setenv("TAHOE_SD_TOP_LEVELS", str(top))
    f = _forest(c)
    info = f.info()
    d = c["depth"]
    if top is not None and (top > d - 1 or (d + 1 - top) % 2):
        assert not info.sd_blocked
        return
    if d < 2 or info.sd_teams < 1:
        assert not info.sd_blocked
        return
    if top is None and not info.sd_blocked:
        assert info.sd_top_levels == d + 1      # small forest: the planner keeps all of it in shared memory
        return
    assert info.sd_blocked and (top is None or info.sd_top_levels == top)
    K = info.sd_top_levels
    assert 1 <= K <= d - 1 and (d + 1 - K) % 2 == 0
    want, wids = oracle.predict_dense(c["nodes"], c["trees"], d, c["data"], float(c["missing"]), want_leaf_ids=True)
    order = f.export(api.EXPORT_TREE_ORDER, np.int32)
    vals, ids = emulate.blocked_traverse(f.export(api.EXPORT_SD_TOP_NODES, np.uint32),
                                         f.export(api.EXPORT_SD_BLOCKS, np.uint32),
                                         f.export(api.EXPORT_SD_ORIG_BOTTOM, np.int32), order, order.shape[0],
                                         d, K, c["trees"], c["data"], float(c["missing"]))
    assert np.array_equal(ids, wids)
    np.testing.assert_allclose(vals.sum(1, dtype=np.float64), want.astype(np.float64), rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("name", ["tiny", "char_w", "short_w", "ties"])
def test_probability_reorder_matches_reference_packer(name):
    """Our swapped trees, tree by tree, equal the reference packer's (Struct.h:1775-1825)."""
    c = load_case(name)
    if "pack_values" not in c:
        pytest.skip("no reference pack in this golden")
    f = _forest(c)
    T, d = c["trees"], c["depth"]
    N = (1 << (d + 1)) - 1
    bl = int(c["pack_bits_length"])
    assert bl == f.info().node_width_bytes
    fid_bits = {1: 5, 2: 13, 4: 29}[bl]
    inter = f.export(api.EXPORT_INTERLEAVED, np.uint32).reshape(N, -1, 2)
    order = f.export(api.EXPORT_TREE_ORDER, np.int32)
    ours = {}
    for slot, t in enumerate(order):
        if t < 0:
            continue
        meta = inter[:, slot, 1]
        ref_style = ((meta & emulate.FIDMASK) & ((1 << fid_bits) - 1)) \
            | np.where(meta & emulate.DEFLEFT, 1 << fid_bits, 0) \
            | np.where(meta & emulate.LEAF, 1 << (fid_bits + 1), 0) \
            | np.where(meta & emulate.EXCH, 1 << (fid_bits + 2), 0)
        ours[inter[:, slot, 0].tobytes() + ref_style.astype(np.uint32).tobytes()] = t
    rv = c["pack_values"].reshape(T, N)
    rb = c["pack_bits"].reshape(T, N).astype(np.uint32)
    seen = set()
    for i in range(T):   # reference tree order is undefined: match as a set
        key = rv[i].view(np.uint32).tobytes() + rb[i].tobytes()
        assert key in ours, f"reference packed tree {i} has no equal in our pack"
        seen.add(ours[key])
    assert len(seen) == len(ours)


def test_wide_forest_has_no_quantized_layout():
    f = _forest(load_case("short_w"))       # 300 columns > 64: one warp tile would not fit
    assert f.info().quantized == 0 and not f.applicable(api.SPLIT_SHARED_FOREST, 100)


@pytest.mark.parametrize("name", ["tiny", "char_w", "ties", "one_feature"])
def test_quantized_layout(oracle, name):
    c = load_case(name)
    f = _forest(c)
    inf = f.info()
    assert inf.quantized == 1 and inf.rank_bits + inf.col_bits + 7 <= 32
    want, wids = oracle.predict_dense(c["nodes"], c["trees"], c["depth"], c["data"],
                                      float(c["missing"]), want_leaf_ids=True)
    table = f.export(api.EXPORT_QUANT_TABLE, np.float32)
    offs = f.export(api.EXPORT_QUANT_OFFSETS, np.int32)
    for col in range(c["cols"]):
        seg = table[offs[col]:offs[col + 1]]
        assert np.all(seg[1:] > seg[:-1])
    shift = 32 - inf.rank_bits
    miss_code = ((1 << inf.rank_bits) - 1) << shift
    assert inf.col_shift >= 7 and inf.rank_bits + inf.col_bits + inf.col_shift <= 32
    col_mask = ((1 << inf.col_bits) - 1) << inf.col_shift
    F = emulate.quantize_rows(c["data"], table, offs, shift, miss_code, float(c["missing"]))
    F2 = emulate.quantize_rows_lut(c["data"], table, offs, f.export(api.EXPORT_QUANT_LUT, np.int32),
                                   f.export(api.EXPORT_QUANT_LUT_MIN, np.float32),
                                   f.export(api.EXPORT_QUANT_LUT_SCALE, np.float32), shift, miss_code,
                                   float(c["missing"]))
    assert np.array_equal(F, F2)          # the bucket index never changes a rank
    vals, ids = emulate.quant_traverse(f.export(api.EXPORT_QUANT_WORDS, np.uint32),
                                       f.export(api.EXPORT_LEAF_ORIGIN, np.int32),
                                       c["trees"], c["depth"], F, col_mask, miss_code, inf.col_shift)
    assert np.array_equal(ids, wids)
    assert np.array_equal(_tree_order_sum(vals).view(np.uint32), want.view(np.uint32))


def test_info_and_chunk_plan():
    c = load_case("susy_like")
    f = _forest(c)
    inf = f.info()
    assert (inf.num_trees, inf.depth, inf.num_cols, inf.is_sparse) == (500, 10, 18, 0)
    assert inf.node_width_bytes == 1 and inf.quantized == 1
    assert inf.num_nodes == 500 * 2047 and abs(inf.mean_depth - 10.0) < 1e-9
    assert inf.trees_per_chunk * inf.num_chunks >= 500 and inf.trees_per_chunk * 8192 < 227 * 1024
    assert not f.applicable(api.SHARED_FOREST, 1000) and f.applicable(api.SPLIT_SHARED_FOREST, 1000)
    assert inf.bytes_quantized == 500 * 2048 * 4


def test_invalid_forests_are_statuses_not_aborts():
    nodes = synth.gen_dense_forest(1, 2, 2, 4)
    with pytest.raises(api.TahoeError):       # fid >= num_cols
        api.Forest.from_dense(nodes, 2, 2, 2)
    bad = nodes.copy()
    bad["bits"] &= 0x7FFFFFFF                  # no leaves at all -> internal node on the bottom level
    with pytest.raises(api.TahoeError):
        api.Forest.from_dense(bad, 2, 2, 4)


def test_sparse_pack_validation():
    nodes = synth.gen_dense_forest(3, 4, 3, 6, leaf_prob=0.3)
    sp, offs, w, ids = synth.dense_to_sparse(nodes, 4, 3)
    f = api.Forest.from_sparse(sp, offs, 6, weights=w, missing=-999.0)
    inf = f.info()
    assert inf.is_sparse == 1 and inf.num_nodes == sp.shape[0] and inf.quantized == 0
    # small trees: every one is whole in shared memory, so both shared-forest strategies apply
    assert f.applicable(api.SPLIT_SHARED_FOREST, 10) and f.applicable(api.SHARED_FOREST, 10)
    assert not f.applicable(api.SHARED_DATA, 10) and f.applicable(api.DIRECT, 10)
    assert inf.sparse_deep_walks == 0.0 and inf.num_chunks == 1
    bad = sp.copy()
    bad["left_idx"][0] = 0
    with pytest.raises(api.TahoeError):
        api.Forest.from_sparse(bad, offs, 6)


def test_predict_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    c = load_case("tiny")
    f = _forest(c)
    out = np.zeros(c["rows"], np.float32)
    with pytest.raises(api.TahoeError):
        f.predict_host(np.ascontiguousarray(c["data"]), out, strategy=api.DIRECT)


@pytest.mark.parametrize("cut", [None, 32, 64])
def test_sparse_breadth_first_layout(oracle, cut, monkeypatch):
    """T8 (sparse forests through shared memory): the breadth-first renumbering walks to the oracle's
    leaves, the shared-memory image is its per-tree prefix, and the planner's deep-walk estimate is sane."""
    if cut is not None:
        monkeypatch.setenv("TAHOE_SPARSE_CUT", str(cut))
    sp, offs, w, lv = synth.gen_sparse_forest(99, 24, 12, 50)
    data = synth.gen_data(98, 600, 50, miss_frac=0.02, nan_frac=0.01)
    want, wids = oracle.predict_sparse(sp, offs, data, -999.0, want_leaf_ids=True)
    f = api.Forest.from_sparse(sp, offs, 50, weights=w, missing=-999.0)
    info = f.info()
    assert info.sparse_cut >= 32 and info.trees_per_chunk >= 1 and info.sparse_fid_bits >= 1
    assert cut is None or info.sparse_cut == cut
    nodes8 = f.export(api.EXPORT_SPARSE_BFS, np.uint32)
    orig = f.export(api.EXPORT_SPARSE_BFS_ORIG, np.int32)
    vals, ids, deep = emulate.sparse_bfs_traverse(nodes8, orig, offs, info.sparse_fid_bits, info.sparse_cut,
                                                  data, -999.0)
    assert np.array_equal(ids, wids)
    np.testing.assert_allclose(vals.sum(1, dtype=np.float64), want.astype(np.float64), rtol=1e-5, atol=1e-5)
    top = f.export(api.EXPORT_SPARSE_TOP, np.uint32).reshape(-1, info.sparse_cut, 2)
    full = nodes8.reshape(-1, 2)
    assert top.shape[0] == info.num_chunks * info.trees_per_chunk >= 24
    fb = info.sparse_fid_bits
    for t in range(24):          # shared-memory form: y = (delta*8) << 17 | fid << 11 | def_left << 1 | leaf
        n = min(int(offs[t + 1] - offs[t]), info.sparse_cut)
        src, img = full[offs[t]:offs[t] + n], top[t, :n]
        leaf = (src[:, 1] & 0x80000000) != 0
        assert np.array_equal(img[:, 0], src[:, 0]) and np.array_equal((img[:, 1] & 1) != 0, leaf)
        assert np.all(img[leaf, 1] == 1)
        y = src[~leaf, 1]
        want_y = ((((y & 0x3FFFFFFF) >> fb) * 8) << 17) | ((y & ((1 << fb) - 1)) << 11) | ((y >> 30) & 1) << 1
        assert np.array_equal(img[~leaf, 1], want_y.astype(np.uint32))
        # the fields the kernel extracts do not overlap
        assert np.array_equal(img[~leaf, 1] >> 17, ((y & 0x3FFFFFFF) >> fb) * 8)
        assert np.array_equal((img[~leaf, 1] & 0x1F800) >> 11, y & ((1 << fb) - 1))
        assert np.all(top[t, n:, 1] == 1) and np.all(top[t, n:, 0] == 0)
    # uniform random features: the measured share of deep walks is near the planner's estimate
    assert abs(deep.mean() - info.sparse_deep_walks) < 0.05


@pytest.mark.parametrize("depth,tpc", [(None, None), (1, None), (3, 6), (6, 12), (12, None)])
def test_sparse_quantised_tops_layout(oracle, depth, tpc, monkeypatch):
    """QT (sparse forests through k_smem_forest_deep): the dense rank-quantised tops plus the two-level
    rank-form blocks below them walk to the oracle's leaves; continuation markers sit exactly where a walk is
    still on an internal node at the top's bottom level; the planner's parked-walk estimate is sane."""
    if depth is not None:
        monkeypatch.setenv("TAHOE_QTOP_DEPTH", str(depth))
    if tpc is not None:                      # several chunks: the markers are chunk-relative
        monkeypatch.setenv("TAHOE_QTOP_TPC", str(tpc))
    sp, offs, w, lv = synth.gen_sparse_forest(99, 24, 12, 50)
    data = synth.gen_data(98, 600, 50, miss_frac=0.02, nan_frac=0.01)
    want, wids = oracle.predict_sparse(sp, offs, data, -999.0, want_leaf_ids=True)
    f = api.Forest.from_sparse(sp, offs, 50, weights=w, missing=-999.0)
    info = f.info()
    assert info.quantized == 0 and info.sparse_top_active == 1
    D = info.sparse_top_depth
    assert 1 <= D <= 12 and (depth is None or D == depth)
    assert info.sparse_top_trees_per_chunk >= 2 and info.sparse_top_trees_per_chunk % info.smem_group == 0
    assert tpc is None or (info.sparse_top_trees_per_chunk <= tpc and info.sparse_top_chunks >= 2)
    assert info.rank_bits + info.col_bits + 7 <= 32 and info.col_shift >= 7
    shift = 32 - info.rank_bits
    miss_code = ((1 << info.rank_bits) - 1) << shift
    col_mask = ((1 << info.col_bits) - 1) << info.col_shift
    table = f.export(api.EXPORT_QUANT_TABLE, np.float32)
    qoffs = f.export(api.EXPORT_QUANT_OFFSETS, np.int32)
    F = emulate.quantize_rows(data, table, qoffs, shift, miss_code, -999.0)
    qwords = f.export(api.EXPORT_QUANT_WORDS, np.uint32)
    assert qwords.size == 24 * (1 << (D + 1))     # (the padding trees of the last group exist on the device only)
    vals, ids, parked, loads = emulate.qtop_traverse(
        qwords, f.export(api.EXPORT_LEAF_ORIGIN, np.int32), f.export(api.EXPORT_SPARSE_DEEP_BLOCKS, np.uint32),
        f.export(api.EXPORT_SPARSE_BLOCK_ORIG, np.int32), f.export(api.EXPORT_SPARSE_BLOCK_OFFSETS, np.int64),
        24, D, F, col_mask, miss_code, info.col_shift, info.sparse_top_trees_per_chunk)
    assert np.array_equal(ids, wids)
    np.testing.assert_allclose(vals.sum(1, dtype=np.float64), want.astype(np.float64), rtol=1e-5, atol=1e-5)
    # a walk is parked exactly when its leaf lies deeper than the top
    leaf_level = lv[offs[:-1][None, :] + wids]
    assert np.array_equal(parked, leaf_level > D)
    # one 32-byte block decides two levels: a walk whose leaf is k >= 1 levels below the top's bottom loads k // 2 + 1 blocks
    assert np.array_equal(loads, np.where(parked, (leaf_level - D) // 2 + 1, 0))
    assert abs(parked.mean() - info.sparse_top_deep_walks) < 0.05
    if D == 12:
        assert info.sparse_top_deep_walks == 0.0 and not parked.any()


def test_sparse_quantised_tops_refused_for_marker_like_leaves():
    """A leaf value whose bits look like a continuation marker (negative quiet NaN) keeps the forest on
    the 8-byte breadth-first tops."""
    sp, offs, w, lv = synth.gen_sparse_forest(7, 4, 6, 10)
    leaf = np.nonzero(sp["bits"] < 0)[0][0]
    sp = sp.copy()
    sp["val"][leaf] = np.array([0xFFC00123], np.uint32).view(np.float32)[0]
    f = api.Forest.from_sparse(sp, offs, 10, weights=w, missing=-999.0)
    info = f.info()
    assert info.sparse_top_depth == 0 and info.sparse_top_active == 0 and info.sparse_cut >= 32
    assert f.applicable(api.SPLIT_SHARED_FOREST, 10)
