"""GPU parity (B200): every strategy through the C ABI against the reference goldens and
the oracle.  Leaf ids bit-exact; raw scores bit-exact for the strategies that sum in tree
order (all thread=sample ones), within 1e-5 relative for SHARED_DATA (lane=tree order)."""
import os

import numpy as np
import pytest

from tahoe_b200 import api, synth
from tests.helpers import (GOLDEN_CASES, TRANSFORMS, bits_equal, golden_preds, l1_mass, load_case,
                           score_close)

pytestmark = pytest.mark.gpu
ORDERED = {api.DIRECT, api.SHARED_DATA_PER_THREAD, api.SHARED_FOREST, api.SPLIT_SHARED_FOREST}


def _run(f, data_np, strategy, want_ids=True):
    import torch
    d = torch.from_numpy(np.ascontiguousarray(data_np)).cuda()
    ids = torch.full((d.shape[0], f.num_trees), -7, dtype=torch.int32, device="cuda") if want_ids else None
    p = f.predict(d, strategy=strategy, leaf_ids=ids)
    torch.cuda.synchronize()
    return p.cpu().numpy(), (ids.cpu().numpy() if want_ids else None)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_all_strategies_match_reference(oracle, name):
    c = load_case(name)
    f = api.Forest.from_dense(c["nodes"], c["trees"], c["depth"], c["cols"], missing=float(c["missing"]))
    want = golden_preds(c, 0, 0.0, 0.0)
    _, wids = oracle.predict_dense(c["nodes"], c["trees"], c["depth"], c["data"],
                                   float(c["missing"]), want_leaf_ids=True)
    mass = l1_mass(c["nodes"], c["trees"], c["depth"], wids)
    ran = 0
    for s in api.ALL_STRATEGIES:
        if not f.applicable(s, c["rows"]):
            with pytest.raises(api.NotApplicable):
                _run(f, c["data"], s)
            continue
        got, ids = _run(f, c["data"], s)
        assert np.array_equal(ids, wids), (name, api.STRATEGY_NAMES[s])
        if s in ORDERED:
            assert bits_equal(got, want), (name, api.STRATEGY_NAMES[s])
        else:
            assert score_close(got, want, mass), (name, api.STRATEGY_NAMES[s])
        got2, _ = _run(f, c["data"], s, want_ids=False)     # perf instantiation, same scores
        assert bits_equal(got, got2)
        ran += 1
    assert ran >= 2
    got, _ = _run(f, c["data"], api.AUTO, want_ids=False)
    assert score_close(got, want, mass)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_shared_data_blocked_walk(oracle, name, monkeypatch):
    """SHARED_DATA through the 32-byte two-level blocks, forced for every legal split of the levels
    (the planner only picks it for forests that outgrow shared memory)."""
    c = load_case(name)
    d = c["depth"]
    _, wids = oracle.predict_dense(c["nodes"], c["trees"], d, c["data"], float(c["missing"]), want_leaf_ids=True)
    want = golden_preds(c, 0, 0.0, 0.0)
    mass = l1_mass(c["nodes"], c["trees"], d, wids)
    for top in range(1, d):
        if (d + 1 - top) % 2:
            continue
        monkeypatch.setenv("TAHOE_SD_TOP_LEVELS", str(top))
        f = api.Forest.from_dense(c["nodes"], c["trees"], d, c["cols"], missing=float(c["missing"]))
        info = f.info()
        if info.sd_teams < 1 or info.sd_top_levels != top:
            continue
        assert info.sd_blocked
        got, ids = _run(f, c["data"], api.SHARED_DATA)
        assert np.array_equal(ids, wids), (name, top)
        assert score_close(got, want, mass), (name, top)


@pytest.mark.parametrize("flags,thr,bias", TRANSFORMS[1:])
def test_output_transforms(oracle, flags, thr, bias):
    c = load_case("char_w")
    f = api.Forest.from_dense(c["nodes"], c["trees"], c["depth"], c["cols"], missing=float(c["missing"]),
                              output=flags, threshold=thr, global_bias=bias)
    want = golden_preds(c, flags, thr, bias)
    for s in api.ALL_STRATEGIES:
        if not f.applicable(s, c["rows"]):
            continue
        got, _ = _run(f, c["data"], s, want_ids=False)
        if flags & 0x100:      # thresholded 0/1: allow flips only where the score sits on the threshold
            raw = oracle.predict_dense(c["nodes"], c["trees"], c["depth"], c["data"], float(c["missing"]),
                                       output=flags & ~0x100, threshold=thr, global_bias=bias)
            bad = got != want
            assert np.all(np.abs(raw[bad] - thr) < 1e-5)
        else:
            assert score_close(got, want), (hex(flags), api.STRATEGY_NAMES[s])


def test_sparse_forest(oracle):
    sp, offs, w, lv = synth.gen_sparse_forest(4242, 64, 12, 50)
    data = synth.gen_data(4343, 3000, 50, miss_frac=0.01, nan_frac=0.005)
    want, wids = oracle.predict_sparse(sp, offs, data, -999.0, want_leaf_ids=True)
    f = api.Forest.from_sparse(sp, offs, 50, weights=w, missing=-999.0)
    import os
    for group in ("4", "1"):      # small-batch kernel (4 walks per thread) / full-machine kernel (free-running lanes)
        os.environ["TAHOE_DIRECT_GROUP"] = group
        try:
            for s in (api.DIRECT, api.SHARED_DATA_PER_THREAD):
                got, ids = _run(f, data, s)
                assert np.array_equal(ids, wids) and bits_equal(got, want), (group, s)
        finally:
            del os.environ["TAHOE_DIRECT_GROUP"]
    with pytest.raises(api.NotApplicable):
        _run(f, data, api.SHARED_DATA)


@pytest.mark.parametrize("cut,tpc,rows", [(None, None, 3000), (32, 7, 3000), (64, 64, 33), (32, 1, 2368 * 32 + 5),
                                          (256, 20, 1)])
def test_sparse_forest_through_shared_memory(oracle, monkeypatch, cut, tpc, rows):
    """k_sparse_smem: tree tops in shared memory, deep walks parked and finished in L2.  Small cuts and
    chunk sizes force many parked walks, partial drains and several chunks; leaf ids stay exact, the
    score is a re-ordered float sum."""
    monkeypatch.setenv("TAHOE_SPARSE_PATH", "t8")      # the quantised tops (QT) are the default path now
    if cut is not None:
        monkeypatch.setenv("TAHOE_SPARSE_CUT", str(cut))
        monkeypatch.setenv("TAHOE_SPARSE_TPC", str(tpc))
    sp, offs, w, lv = synth.gen_sparse_forest(4242, 64, 12, 50)
    data = synth.gen_data(4343, rows, 50, miss_frac=0.01, nan_frac=0.005)
    want, wids = oracle.predict_sparse(sp, offs, data, -999.0, want_leaf_ids=True)
    f = api.Forest.from_sparse(sp, offs, 50, weights=w, missing=-999.0)
    info = f.info()
    if cut is not None:
        assert info.sparse_cut == cut and info.trees_per_chunk == tpc and info.num_chunks == -(-64 // tpc)
    assert f.applicable(api.SPLIT_SHARED_FOREST, rows)
    leafvals = np.abs(sp["val"][offs[:-1][None, :] + wids]).sum(1)       # L1 mass of each row's summands
    got, ids = _run(f, data, api.SPLIT_SHARED_FOREST)
    assert np.array_equal(ids, wids)
    assert score_close(got, want, leafvals)
    # the perf instantiation (other number of tree sequences per lane: another order again)
    got2, _ = _run(f, data, api.SPLIT_SHARED_FOREST, want_ids=False)
    assert score_close(got2, want, leafvals)
    got2b, _ = _run(f, data, api.SPLIT_SHARED_FOREST, want_ids=False)
    assert bits_equal(got2, got2b)                                        # reproducible run to run
    if f.applicable(api.SHARED_FOREST, rows):
        got3, _ = _run(f, data, api.SHARED_FOREST, want_ids=False)
        assert bits_equal(got2, got3)


@pytest.mark.parametrize("depth,tpc,plan,rows,cols", [
    (None, None, None, 3000, 50), (2, 6, "16,1,6", 3000, 18), (5, 12, "8,1,6", 33, 50), (1, 2, "8,1,2", 2368 * 32 + 5, 50),
    (8, 64, "4,1,4", 1, 50), (12, None, "8,1,2", 700, 50), (3, 4, "16,1,4", 40_000, 24), (4, 6, "16,1,6", 5000, 28),
    (2, 6, "4,1,6", 3000, 64), (6, 12, "8,1,4", 148 * 8 * 32 * 3 + 17, 50),
    # 16 warps on rows wider than 32 columns: pairs of warps share a tile
    (3, 12, "16,1,6", 148 * 8 * 32 * 2 + 33, 50), (2, 4, "16,1,2", 5000, 64), (5, 24, "16,1,4", 31, 40), (7, None, "16,1,6", 70_000, 50)])
def test_sparse_forest_quantised_tops(oracle, monkeypatch, depth, tpc, plan, rows, cols):
    """k_smem_forest_deep on the QT layout: dense rank-quantised tops in shared memory, walks still on an
    internal node below them parked in lane-private rings and advanced one level per traversal group in
    L2, across the tile boundary (two tile buffers per warp).  Shallow tops and small chunks force many
    parked walks, full rings, leftovers at tile and chunk ends, several tiles per warp and several
    chunks; 16 / 8 / 4 warps, 32- and 64-column register staging, pairs of warps sharing a tile; leaf ids
    stay exact, the score is a re-ordered float sum."""
    if depth is not None:
        monkeypatch.setenv("TAHOE_QTOP_DEPTH", str(depth))
    if tpc is not None:
        monkeypatch.setenv("TAHOE_QTOP_TPC", str(tpc))
    if plan is not None:
        monkeypatch.setenv("TAHOE_SMEM_PLAN", plan)
    # (16 warps: the parking lists leave room for two TMA buffers per warp only with narrow rows)
    sp, offs, w, lv = synth.gen_sparse_forest(4242, 64, 12, cols)
    data = synth.gen_data(4343, rows, cols, miss_frac=0.01, nan_frac=0.005)
    want, wids = oracle.predict_sparse(sp, offs, data, -999.0, want_leaf_ids=True)
    f = api.Forest.from_sparse(sp, offs, cols, weights=w, missing=-999.0)
    info = f.info()
    assert info.sparse_top_active == 1
    if depth is not None:
        assert info.sparse_top_depth == depth
    if plan is not None:
        assert (info.smem_warps, info.smem_staged, info.smem_group) == tuple(int(x) for x in plan.split(","))
    if tpc is not None:
        assert info.sparse_top_trees_per_chunk <= max(tpc, info.smem_group)
        assert info.sparse_top_chunks > 1 or tpc >= 64
    assert f.applicable(api.SPLIT_SHARED_FOREST, rows)
    leafvals = np.abs(sp["val"][offs[:-1][None, :] + wids]).sum(1)       # L1 mass of each row's summands
    got, ids = _run(f, data, api.SPLIT_SHARED_FOREST)
    assert np.array_equal(ids, wids)
    assert score_close(got, want, leafvals)
    got2, _ = _run(f, data, api.SPLIT_SHARED_FOREST, want_ids=False)      # the perf instantiation
    assert score_close(got2, want, leafvals)
    got2b, _ = _run(f, data, api.SPLIT_SHARED_FOREST, want_ids=False)
    assert bits_equal(got2, got2b)                                        # reproducible run to run
    if depth == 12:                                                        # whole trees in the tops: tree order, CPU bits
        assert info.sparse_top_deep_walks == 0.0 and bits_equal(got2, want)
    if f.applicable(api.SHARED_FOREST, rows):
        got3, _ = _run(f, data, api.SHARED_FOREST, want_ids=False)
        assert bits_equal(got2, got3)


@pytest.mark.parametrize("rows", [0, 1, 31, 32, 33, 2368 * 32 + 5])
def test_ragged_batches(oracle, rows):
    c = load_case("char_w")
    f = api.Forest.from_dense(c["nodes"], c["trees"], c["depth"], c["cols"], missing=float(c["missing"]))
    data = synth.gen_data(555, rows, c["cols"], miss_frac=0.01) if rows else np.zeros((0, c["cols"]), np.float32)
    want = oracle.predict_dense(c["nodes"], c["trees"], c["depth"], data, float(c["missing"])) if rows else np.zeros(0, np.float32)
    for s in api.ALL_STRATEGIES:
        if f.applicable(s, rows):
            got, _ = _run(f, data, s, want_ids=False)
            assert got.shape == (rows,)
            assert (bits_equal if s in ORDERED else score_close)(got, want), (rows, s)


def test_multi_chunk_split_and_host_path(oracle):
    """C2-shaped forest (500 x depth 10 x 18 cols): many shared-memory chunks; device generator
    equals the host generator; tahoe_predict_host equals tahoe_predict."""
    import torch
    c = load_case("susy_like")
    f = api.Forest.from_dense(c["nodes"], c["trees"], c["depth"], c["cols"], missing=-999.0)
    assert f.info().num_chunks > 1
    rows = 100_003
    d = api.gen_data_device(777, rows, 18, miss_frac=0.001)
    host = synth.gen_data(777, rows, 18, miss_frac=0.001)
    assert np.array_equal(d.cpu().numpy().view(np.uint32), host.view(np.uint32))
    p = f.predict(d, strategy=api.SPLIT_SHARED_FOREST)
    torch.cuda.synchronize()
    assert f.launches() == 2 or os.environ.get("TAHOE_TOP_TREES") or os.environ.get("TAHOE_QUANT_TILES")   # (forced variants: more launches)
    sub = np.r_[0:4096, rows - 4096:rows]
    want = oracle.predict_dense(c["nodes"], 500, 10, host[sub], -999.0)
    assert bits_equal(p.cpu().numpy()[sub], want)
    p2 = f.predict(d, strategy=api.DIRECT)
    assert bits_equal(p2.cpu().numpy(), p.cpu().numpy())        # strategy vs strategy, all rows
    p3 = f.predict(d, strategy=api.SHARED_DATA)        # lane = tree: another summation order
    _, ids_all = oracle.predict_dense(c["nodes"], 500, 10, host, -999.0, want_leaf_ids=True)
    assert score_close(p3.cpu().numpy(), p.cpu().numpy(), l1_mass(c["nodes"], 500, 10, ids_all))
    ph = f.predict_host(host, strategy=api.SPLIT_SHARED_FOREST)
    assert bits_equal(ph, p.cpu().numpy())
    assert bits_equal(golden_preds(c, 0, 0.0, 0.0),
                      f.predict_host(np.ascontiguousarray(c["data"]), strategy=api.SPLIT_SHARED_FOREST))


@pytest.mark.parametrize("cluster,trees,rows", [(2, 40, 200_003), (4, 36, 50_001), (2, 12, 33), (4, 44, 1)])
def test_cluster_pipeline_is_bit_exact(oracle, monkeypatch, cluster, trees, rows):
    """k_smem_forest_cluster (kernels_smem_cluster.cuh): S CTAs hold S consecutive chunks, tiles arrive by TMA
    multicast, partial sums pass CTA -> CTA through distributed shared memory IN TREE ORDER -- the scores must be
    the CPU's bits.  Forced on a Higgs-shaped forest (depth 12: one group of 4 trees per chunk); chunk counts that
    do and do not divide by S, missing values, one tile, one row."""
    import torch
    d, C = 12, 28
    monkeypatch.setenv("TAHOE_SMEM_PLAN", "12,0,4")
    monkeypatch.setenv("TAHOE_CLUSTER", str(cluster))
    nodes = synth.gen_dense_forest(4100 + trees, trees, d, C, leaf_prob=0.02)
    f = api.Forest.from_dense(nodes, trees, d, C, missing=-999.0)
    inf = f.info()
    assert (inf.smem_warps, inf.smem_staged, inf.smem_group, inf.trees_per_chunk) == (12, 0, 4, 4)
    assert inf.num_chunks == trees // 4 and inf.num_chunks >= 2 * cluster or rows <= 33
    x = api.gen_data_device(4200 + trees, rows, C, miss_frac=0.002)
    got = f.predict(x, strategy=api.SPLIT_SHARED_FOREST).cpu().numpy()
    ref = f.predict(x, strategy=api.DIRECT).cpu().numpy()
    assert bits_equal(got, ref)
    n = min(rows, 2048)
    host = synth.gen_data(4200 + trees, n, C, miss_frac=0.002)
    assert bits_equal(got[:n], oracle.predict_dense(nodes, trees, d, host, -999.0))
    got2 = f.predict(x, strategy=api.SPLIT_SHARED_FOREST).cpu().numpy()      # again: barriers and phases end clean
    assert bits_equal(got2, ref)
    monkeypatch.setenv("TAHOE_CLUSTER", "0")
    f0 = api.Forest.from_dense(nodes, trees, d, C, missing=-999.0)
    assert bits_equal(f0.predict(x, strategy=api.SPLIT_SHARED_FOREST).cpu().numpy(), ref)


def test_two_streams_share_the_handle_scratch_safely():
    """tahoe_predict on two streams without a host sync in between: both use the handle's one workspace (rank tiles,
    tile flags), so the second call must wait for the first ON THE DEVICE (the event in predict_on_ws0)."""
    import torch
    c = load_case("susy_like")
    f = api.Forest.from_dense(c["nodes"], c["trees"], c["depth"], c["cols"], missing=-999.0)
    xa = api.gen_data_device(91, 300_000, 18, miss_frac=0.001)
    xb = api.gen_data_device(92, 200_000, 18)
    want_a = f.predict(xa, strategy=api.SPLIT_SHARED_FOREST).clone()
    want_b = f.predict(xb, strategy=api.SPLIT_SHARED_FOREST).clone()
    torch.cuda.synchronize()
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    pa, pb = torch.empty_like(want_a), torch.empty_like(want_b)
    for _ in range(3):
        f.predict(xa, pa, strategy=api.SPLIT_SHARED_FOREST, stream=sa)
        f.predict(xb, pb, strategy=api.SPLIT_SHARED_FOREST, stream=sb)     # no sync: overlaps the first on the device unless ordered
    torch.cuda.synchronize()
    assert torch.equal(pa, want_a) and torch.equal(pb, want_b)


def test_measure_peaks_and_copy_baseline():
    import torch
    pk = api.measure_peaks()
    assert pk["sm_count"] > 0 and pk["l2_bytes"] > 0
    assert pk["smem_read_gbs"] > pk["l2_read_gbs"] > 1000.0 and pk["l2_sector_gathers_per_s"] > 1e10
    c = load_case("susy_like")
    f = api.Forest.from_dense(c["nodes"], c["trees"], c["depth"], c["cols"], missing=-999.0)
    host = synth.gen_data(93, 50_001, 18)
    want = f.predict_host(host, strategy=api.SPLIT_SHARED_FOREST)
    scratch = np.empty(50_001, dtype=np.float32)
    f.predict_host(host, scratch, strategy=api.SPLIT_SHARED_FOREST, copy_only=True)     # same slabs, no kernels: must not fail
    assert f.launches() == 0
    assert bits_equal(f.predict_host(host, strategy=api.SPLIT_SHARED_FOREST), want)
    with pytest.raises(ValueError):
        f.predict_host(host.astype(np.float64), strategy=api.SPLIT_SHARED_FOREST)     # the C ABI takes bare pointers: checked here
    with pytest.raises(ValueError):
        f.predict(torch.zeros((10, 17), device="cuda"))


@pytest.mark.parametrize("tiles", ["0", "1"])
def test_quantise_from_rows_and_from_float_tiles_agree(oracle, monkeypatch, tiles):
    """k_quantize reads the row-major matrix directly, or (strided threshold tables: configs 3 and 4) the column-major
    float tiles k_float_tiles wrote into the same buffer, ranks going back in place.  Both forms, forced on a dense and
    on a sparse forest with missing values, NaN and ragged tile ends: the CPU's bits / leaf ids."""
    import torch
    monkeypatch.setenv("TAHOE_QUANT_TILES", tiles)
    c = load_case("susy_like")
    f = api.Forest.from_dense(c["nodes"], c["trees"], c["depth"], c["cols"], missing=-999.0)
    for rows in (1, 33, 70_001):
        host = synth.gen_data(55, rows, 18, miss_frac=0.01, nan_frac=0.01)
        x = torch.from_numpy(host).cuda()
        got = f.predict(x, strategy=api.SPLIT_SHARED_FOREST).cpu().numpy()
        assert f.launches() == (3 if tiles == "1" else 2)
        assert bits_equal(got, oracle.predict_dense(c["nodes"], c["trees"], c["depth"], host, -999.0)), rows
    sp, offs, w, _ = synth.gen_sparse_forest(56, 48, 14, 50)
    fs = api.Forest.from_sparse(sp, offs, 50, weights=w, missing=-999.0)
    host = synth.gen_data(57, 9_999, 50, miss_frac=0.01, nan_frac=0.01)
    want, wids = oracle.predict_sparse(sp, offs, host, -999.0, want_leaf_ids=True)
    ids = torch.zeros((host.shape[0], 48), dtype=torch.int32, device="cuda")
    got = fs.predict(torch.from_numpy(host).cuda(), strategy=api.SPLIT_SHARED_FOREST, leaf_ids=ids).cpu().numpy()
    assert np.array_equal(ids.cpu().numpy(), wids)
    mass = np.abs(sp["val"][offs[:-1][None, :] + wids]).sum(1)
    assert score_close(got, want, mass)


@pytest.mark.parametrize("plan,trees,depth,cols,rows", [
    ("16,2,5", 203, 10, 18, 148 * 16 * 32 + 77), ("16,0,5", 203, 10, 18, 5000), ("16,1,5", 203, 10, 18, 5000),
    ("16,2,5", 47, 12, 28, 148 * 16 * 32 * 2 + 5), ("12,2,5", 47, 12, 28, 33), ("8,2,3", 10, 11, 28, 1),
    ("16,2,2", 7, 6, 40, 70_001), ("4,2,1", 5, 3, 64, 2049), ("16,2,6", 600, 8, 5, 300_000)])
def test_smem_tile_modes_are_bit_exact(oracle, monkeypatch, plan, trees, depth, cols, rows):
    """k_smem_forest under every way of filling a warp's tile buffer -- two buffers by TMA (0), one + registers (1),
    SINGLE (2: one buffer by TMA, refilled from L2 after the walk) -- forced by TAHOE_SMEM_PLAN on dense forests of
    several depths: scores are the CPU's bits, leaf ids the oracle's, twice in a row (mbarrier phases end clean);
    several chunks, one tile, one row, rows with missing values and NaN."""
    import torch
    monkeypatch.setenv("TAHOE_SMEM_PLAN", plan)
    nodes = synth.gen_dense_forest(5200 + trees, trees, depth, cols, leaf_prob=0.03)
    f = api.Forest.from_dense(nodes, trees, depth, cols, missing=-999.0)
    inf = f.info()
    if not f.applicable(api.SPLIT_SHARED_FOREST, rows):
        pytest.skip("plan does not fit this forest")
    assert (inf.smem_warps, inf.smem_staged, inf.smem_group) == tuple(int(x) for x in plan.split(","))
    x = api.gen_data_device(5300 + trees, rows, cols, miss_frac=0.004)
    ref = f.predict(x, strategy=api.DIRECT).cpu().numpy()
    for _ in range(2):
        assert bits_equal(f.predict(x, strategy=api.SPLIT_SHARED_FOREST).cpu().numpy(), ref)
    n = min(rows, 1500)
    host = synth.gen_data(5300 + trees, n, cols, miss_frac=0.004, nan_frac=0.002)
    want, wids = oracle.predict_dense(nodes, trees, depth, host, -999.0, want_leaf_ids=True)
    got, ids = _run(f, host, api.SPLIT_SHARED_FOREST)
    assert np.array_equal(ids, wids) and bits_equal(got, want)
