"""GPU: seeded random shapes through every applicable strategy against the oracle -- forests and
batches nobody hand-picked (odd column counts, one tree, one row, shallow and deep, mostly-missing
rows, every transform).  Leaf ids bit-exact; scores within 1e-5 of the L1 mass of the summands."""
import numpy as np
import pytest

from tahoe_b200 import api, synth
from tests.helpers import score_close

import os

pytestmark = pytest.mark.gpu
SOAK = int(os.environ.get("TAHOE_FUZZ_SOAK", "1"))      # x the seed counts (soak runs: TAHOE_FUZZ_SOAK=10)


def _run(f, data, s, ids=True):
    import torch
    d = torch.from_numpy(np.ascontiguousarray(data)).cuda()
    out = torch.full((d.shape[0], f.num_trees), -7, dtype=torch.int32, device="cuda") if ids else None
    p = f.predict(d, strategy=s, leaf_ids=out)
    torch.cuda.synchronize()
    return p.cpu().numpy(), (out.cpu().numpy() if ids else None)


def _check_all(f, data, want, wids, mass, tag):
    ran = 0
    for s in api.ALL_STRATEGIES:
        if not f.applicable(s, data.shape[0]):
            continue
        got, ids = _run(f, data, s)
        assert np.array_equal(ids, wids), (tag, api.STRATEGY_NAMES[s])
        assert score_close(got, want, mass), (tag, api.STRATEGY_NAMES[s])
        got2, _ = _run(f, data, s, ids=False)
        assert score_close(got2, want, mass), (tag, api.STRATEGY_NAMES[s])
        ran += 1
    assert ran >= 1, tag


@pytest.mark.parametrize("seed", range(20 * SOAK))
def test_random_sparse_forests(oracle, seed):
    rng = np.random.default_rng(1000 + seed)
    T = int(rng.choice([1, 2, 5, 6, 7, 31, 64, 130]))
    depth = int(rng.integers(1, 17))
    cols = int(rng.choice([1, 2, 17, 32, 33, 50, 64]))
    rows = int(rng.choice([1, 31, 33, 1000, 4097, 148 * 16 * 32 + 1]))
    leaf_prob = float(rng.choice([0.1, 0.3, 0.6]))
    miss = float(rng.choice([0.0, 0.02, 0.5]))
    sp, offs, w, lv = synth.gen_sparse_forest(500 + seed, T, depth, cols, leaf_prob=leaf_prob, spine=bool(seed % 2))
    data = synth.gen_data(600 + seed, rows, cols, miss_frac=miss, nan_frac=0.01)
    output = int(rng.choice([0, 0x1, 0x10, 0x11]))
    want, wids = oracle.predict_sparse(sp, offs, data, -999.0, output=output, global_bias=0.25, want_leaf_ids=True)
    f = api.Forest.from_sparse(sp, offs, cols, weights=w, missing=-999.0, output=output, global_bias=0.25)
    mass = np.abs(sp["val"][offs[:-1][None, :] + wids]).sum(1) + 0.25
    _check_all(f, data, want, wids, mass, (seed, T, depth, cols, rows, leaf_prob, miss, output))


@pytest.mark.parametrize("seed", range(8 * SOAK))
def test_random_dense_forests(oracle, seed):
    rng = np.random.default_rng(2000 + seed)
    T = int(rng.choice([1, 3, 32, 33, 100, 257]))
    depth = int(rng.integers(1, 14))          # 10+: deep plans (one traversal group per chunk: the cluster pipeline)
    cols = int(rng.choice([1, 3, 18, 28, 64, 65, 200]))
    rows = int(rng.choice([1, 32, 77, 5000, 148 * 16 * 32 + 31]))
    nodes = synth.gen_dense_forest(700 + seed, T, depth, cols, leaf_prob=float(rng.choice([0.0, 0.05, 0.3])))
    data = synth.gen_data(800 + seed, rows, cols, miss_frac=float(rng.choice([0.0, 0.05])), nan_frac=0.01)
    want, wids = oracle.predict_dense(nodes, T, depth, data, -999.0, want_leaf_ids=True)
    f = api.Forest.from_dense(nodes, T, depth, cols, missing=-999.0)
    npt = (1 << (depth + 1)) - 1
    mass = np.abs(nodes["val"].reshape(T, npt)[np.arange(T)[None, :], wids]).sum(1)
    _check_all(f, data, want, wids, mass, (seed, T, depth, cols, rows))


@pytest.mark.parametrize("seed", range(6 * SOAK))
def test_random_deep_sparse_forests(oracle, seed, monkeypatch):
    """Bigger, deeper sparse forests with random top depths and chunk sizes forced on the planner:
    many deep walks per row, full rings, several tiles per warp, several chunks, pairs of warps."""
    rng = np.random.default_rng(3000 + seed)
    T = int(rng.choice([100, 200, 500]))
    depth = int(rng.integers(10, 21))
    cols = int(rng.choice([20, 33, 50, 64]))
    rows = int(rng.choice([20_000, 60_001, 148 * 16 * 32 * 2 + 7]))
    leaf_prob = float(rng.choice([0.2, 0.3, 0.45]))
    if rng.random() < 0.7:
        monkeypatch.setenv("TAHOE_QTOP_DEPTH", str(int(rng.integers(1, 9))))
    if rng.random() < 0.5:
        monkeypatch.setenv("TAHOE_QTOP_TPC", str(int(rng.choice([6, 12, 24, 48]))))
    sp, offs, w, lv = synth.gen_sparse_forest(900 + seed, T, depth, cols, leaf_prob=leaf_prob)
    data = synth.gen_data(950 + seed, rows, cols, miss_frac=float(rng.choice([0.0, 0.01, 0.2])), nan_frac=0.002)
    want, wids = oracle.predict_sparse(sp, offs, data, -999.0, want_leaf_ids=True)
    f = api.Forest.from_sparse(sp, offs, cols, weights=w, missing=-999.0)
    # (a forced top depth with a giant chunk can push a block index past the marker's 22 bits -- soak seed 45: 200 trees to
    # level 20, 10.8 M nodes in ONE chunk at D = 2 -- and the packer then keeps the forest on the 8-byte tops: legal, and
    # the checks below hold for that path too)
    inf = f.info()
    assert inf.sparse_top_active == 1 or (inf.sparse_cut >= 1 and f.applicable(api.SPLIT_SHARED_FOREST, rows))
    mass = np.abs(sp["val"][offs[:-1][None, :] + wids]).sum(1)
    for s in (api.SPLIT_SHARED_FOREST, api.DIRECT):
        got, ids = _run(f, data, s)
        assert np.array_equal(ids, wids), (seed, api.STRATEGY_NAMES[s])
        assert score_close(got, want, mass), (seed, api.STRATEGY_NAMES[s])
        got2, _ = _run(f, data, s, ids=False)
        assert score_close(got2, want, mass), (seed, api.STRATEGY_NAMES[s])
        got3, _ = _run(f, data, s, ids=False)
        assert np.array_equal(got2.view(np.uint32), got3.view(np.uint32))     # reproducible run to run
    # host buffers in, host scores out: wave-aligned slabs through the same kernels
    got_h = f.predict_host(data, strategy=api.SPLIT_SHARED_FOREST)
    assert score_close(got_h, want, mass), seed
