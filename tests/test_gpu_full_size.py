"""GPU parity at BASELINE.json's FULL sizes (configs 1-4), through the C ABI.

The oracle is single-digit microseconds per (row, tree) visit chain, so at these sizes it checks a
seeded row subset (first rows, last rows, a stride through the middle) bit for bit; every row is then
covered by size-independent properties: all strategies that sum in tree order must agree bit for bit
with each other on ALL rows, SHARED_DATA within the north-star tolerance, and the host-buffer path
with the device path."""
import numpy as np
import pytest

from tahoe_b200 import api, synth
from tests.helpers import abs_leaf_forest, bits_equal, l1_mass, l1_mass_sparse, score_close, score_tight

pytestmark = pytest.mark.gpu
ORDERED = (api.DIRECT, api.SHARED_DATA_PER_THREAD, api.SHARED_FOREST, api.SPLIT_SHARED_FOREST)


def _subset(rows, k=2048):
    mid = np.arange(0, rows, max(1, rows // k))[:k]
    return np.unique(np.r_[0:min(k, rows), mid, max(0, rows - k):rows])


def _device_rows(cfg, rows, miss_frac):
    c = synth.CONFIGS[cfg]
    return api.gen_data_device(synth.DATA_SEED[cfg], rows, c["cols"], miss_frac=miss_frac)


def _host_rows(cfg, idx, miss_frac):
    """the same generator on the host, only for the rows in idx (element index = row*cols + col)"""
    c = synth.CONFIGS[cfg]
    out = np.empty((idx.shape[0], c["cols"]), np.float32)
    for i, r in enumerate(idx):
        out[i] = synth.gen_data(synth.DATA_SEED[cfg], 1, c["cols"], miss_frac=miss_frac, row0=int(r))[0]
    return out


@pytest.mark.parametrize("cfg,miss_frac", [("C1", 0.0), ("C2", 0.001), ("C3", 0.0)])
def test_dense_config_full_size(oracle, cfg, miss_frac):
    import torch
    c = synth.CONFIGS[cfg]
    R, T, d, C = c["rows"], c["trees"], c["depth"], c["cols"]
    nodes = synth.gen_dense_forest(synth.FOREST_SEED[cfg], T, d, C)
    f = api.Forest.from_dense(nodes, T, d, C, missing=-999.0)
    data = _device_rows(cfg, R, miss_frac)
    idx = _subset(R)
    host = _host_rows(cfg, idx, miss_frac)
    assert np.array_equal(data[torch.from_numpy(idx).cuda()].cpu().numpy().view(np.uint32), host.view(np.uint32))
    want, wids = oracle.predict_dense(nodes, T, d, host, -999.0, want_leaf_ids=True)
    ref = None
    ran = []
    for s in api.ALL_STRATEGIES:
        if not f.applicable(s, R):
            continue
        got = f.predict(data, strategy=s)
        torch.cuda.synchronize()
        g = got.cpu().numpy()
        if s in ORDERED:
            assert bits_equal(g[idx], want), (cfg, api.STRATEGY_NAMES[s])          # oracle, subset
            if ref is None:
                ref = g
            else:
                assert bits_equal(g, ref), (cfg, api.STRATEGY_NAMES[s])            # strategy vs strategy, ALL rows
        else:
            mass = l1_mass(nodes, T, d, wids)
            assert score_close(g[idx], want, mass), (cfg, api.STRATEGY_NAMES[s])
            assert score_tight(g[idx], want, mass), (cfg, api.STRATEGY_NAMES[s])
            if ref is not None:
                # every row, against the tree-order sum, bounded by that row's OWN L1 mass: the score of the
                # same trees with |leaf value| leaves, computed on the device by an in-order strategy
                fa = api.Forest.from_dense(abs_leaf_forest(nodes), T, d, C, missing=-999.0)
                in_order = api.SPLIT_SHARED_FOREST if fa.applicable(api.SPLIT_SHARED_FOREST, R) else api.DIRECT
                mass_all = fa.predict(data, strategy=in_order).cpu().numpy()
                fa.close()
                assert np.allclose(mass_all[idx], mass, rtol=1e-5)
                assert score_close(g, ref, mass_all) and score_tight(g, ref, mass_all), (cfg, api.STRATEGY_NAMES[s])
        ran.append(s)
    assert len(ran) >= 2
    # leaf ids on the subset rows, through the strategy the model picks for the full batch
    pick, _ = f.select(R)
    sub = data[torch.from_numpy(idx).cuda()].contiguous()
    ids = torch.zeros((idx.shape[0], T), dtype=torch.int32, device="cuda")
    f.predict(sub, strategy=pick, leaf_ids=ids)
    torch.cuda.synchronize()
    assert np.array_equal(ids.cpu().numpy(), wids)
    if cfg in ("C2", "C3"):
        # host-buffer path = device path on every row, and REPEATEDLY: from the second call on nothing serialises the
        # slab streams any more (no workspace growth), so the next slab's kernels run beside the traversal of the one
        # before -- the regime in which the cluster pipeline returned wrong tiles (profiles/r02_cluster_pipeline.md)
        dev = f.predict(data, strategy=pick).cpu().numpy()
        h_data = torch.empty(tuple(data.shape), dtype=torch.float32, pin_memory=True)
        h_data.copy_(data)
        h_preds = torch.empty(R, dtype=torch.float32, pin_memory=True)
        for _ in range(3):
            f.predict_host(h_data, h_preds, strategy=pick)
            assert bits_equal(h_preds.numpy(), dev), cfg


def test_sparse_config_full_size(oracle):
    """config 4: 2000 deep unbalanced trees (<= 20 levels), 1M rows; the dense layout cannot hold it."""
    import torch
    c = synth.CONFIGS["C4"]
    R = c["rows"]
    sp, offs, w, lv = synth.gen_sparse_forest(synth.FOREST_SEED["C4"], c["trees"], c["depth"], c["cols"])
    assert int(lv.max()) == 20 and sp.shape[0] > 5_000_000
    f = api.Forest.from_sparse(sp, offs, c["cols"], weights=w, missing=-999.0)
    data = _device_rows("C4", R, 0.001)
    idx = _subset(R, 1024)
    host = _host_rows("C4", idx, 0.001)
    want, wids = oracle.predict_sparse(sp, offs, host, -999.0, want_leaf_ids=True)
    a = f.predict(data, strategy=api.DIRECT).cpu().numpy()
    b = f.predict(data, strategy=api.SHARED_DATA_PER_THREAD).cpu().numpy()
    assert bits_equal(a[idx], want) and bits_equal(a, b)
    sub = data[torch.from_numpy(idx).cuda()].contiguous()
    ids = torch.zeros((idx.shape[0], c["trees"]), dtype=torch.int32, device="cuda")
    f.predict(sub, strategy=api.DIRECT, leaf_ids=ids)
    torch.cuda.synchronize()
    assert np.array_equal(ids.cpu().numpy(), wids)
    # quantised tree tops in shared memory (QT), deep walks finished in L2: exact leaves, re-ordered float sum
    info = f.info()
    assert f.applicable(api.SPLIT_SHARED_FOREST, R) and info.sparse_top_active == 1
    assert info.sparse_top_chunks > 1 and info.sparse_top_deep_walks > 0
    s = f.predict(data, strategy=api.SPLIT_SHARED_FOREST).cpu().numpy()
    # scores against the ORACLE on the subset, bounded by each row's real L1 mass (~1000 here) ...
    mass = l1_mass_sparse(sp, offs, wids)
    assert score_close(s[idx], want, mass) and score_tight(s[idx], want, mass)
    # ... and against the tree-order sum (DIRECT == oracle bits) on ALL rows, each row bounded by its own mass:
    # the same trees with |leaf value| leaves, summed in order on the device
    fa = api.Forest.from_sparse(abs_leaf_forest(sp), offs, c["cols"], weights=w, missing=-999.0)
    mass_all = fa.predict(data, strategy=api.DIRECT).cpu().numpy()
    fa.close()
    assert np.allclose(mass_all[idx], mass, rtol=1e-5)
    assert score_close(s, a, mass_all) and score_tight(s, a, mass_all)
    ids.zero_()
    f.predict(sub, strategy=api.SPLIT_SHARED_FOREST, leaf_ids=ids)
    torch.cuda.synchronize()
    assert np.array_equal(ids.cpu().numpy(), wids)
