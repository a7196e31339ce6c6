"""Importers (tahoe_b200/importers.py): trained scikit-learn ensembles and XGBoost
JSON models become Tahoe compact forests that select the SAME leaves as the trainer's own predict --
checked with the oracle on CPU and through the C ABI on the GPU."""
import json
import math

import numpy as np
import pytest

from tahoe_b200 import importers

sk = pytest.importorskip("sklearn.ensemble")


def _data(rows=600, cols=12, seed=5):
    rng = np.random.default_rng(seed)
    X = rng.uniform(-1, 1, size=(rows, cols)).astype(np.float32)
    y = (np.sin(3 * X[:, 0]) + X[:, 1] * X[:, 2] + 0.3 * (X[:, 3] > 0.2) + 0.05 * rng.normal(size=rows)).astype(np.float64)
    return X, y


def _with_ties(model_trees, X):
    """put exact (float32-rounded) thresholds of the trained trees into the data: ties must fall the trainer's way"""
    X = X.copy()
    k = 0
    for tr in model_trees:
        t = tr.tree_
        for n in np.nonzero(t.children_left >= 0)[0][:3]:
            X[k % X.shape[0], t.feature[n]] = np.float32(t.threshold[n])
            k += 1
    return X


def _leaf_ids_as_source(imp, lids):
    return np.stack([imp.source_ids[t][lids[:, t]] for t in range(imp.num_trees)], axis=1)


def _l1(imp, lids):
    return np.abs(imp.nodes["val"][imp.tree_offsets[:-1][None, :] + lids]).sum(1) + abs(imp.global_bias)


@pytest.mark.parametrize("kind", ["rf", "extra", "gbr", "tree"])
def test_sklearn_regressors_match_their_own_predict(oracle, kind):
    X, y = _data()
    if kind == "rf":
        m = sk.RandomForestRegressor(n_estimators=25, max_depth=7, random_state=1).fit(X, y)
        trees = list(m.estimators_)
    elif kind == "extra":
        m = sk.ExtraTreesRegressor(n_estimators=15, max_depth=9, random_state=2).fit(X, y)
        trees = list(m.estimators_)
    elif kind == "gbr":
        m = sk.GradientBoostingRegressor(n_estimators=60, max_depth=3, learning_rate=0.1, random_state=3).fit(X, y)
        trees = [e[0] for e in m.estimators_]
    else:
        from sklearn.tree import DecisionTreeRegressor
        m = DecisionTreeRegressor(max_depth=10, random_state=4).fit(X, y)
        trees = [m]
    Xt = _with_ties(trees, X)
    imp = importers.from_sklearn(m)
    assert imp.num_trees == len(trees) and imp.num_cols == X.shape[1]
    got, lids = oracle.predict_sparse(imp.nodes, imp.tree_offsets, Xt, -np.inf, output=imp.output,
                                      global_bias=imp.global_bias, want_leaf_ids=True)
    want_leaves = np.stack([tr.apply(Xt) for tr in trees], axis=1)
    assert np.array_equal(_leaf_ids_as_source(imp, lids), want_leaves)      # same leaves, ties included
    want = m.predict(Xt)
    scale = _l1(imp, lids) / (imp.num_trees if imp.output & importers.OUT_AVG else 1)
    assert np.all(np.abs(got - want) <= 1e-5 * np.maximum(scale, 1.0))


def test_sklearn_binary_gradient_boosting_is_a_sigmoid(oracle):
    X, y = _data(seed=9)
    m = sk.GradientBoostingClassifier(n_estimators=40, max_depth=3, random_state=0).fit(X, (y > 0.2).astype(int))
    imp = importers.from_sklearn(m)
    assert imp.output == importers.OUT_SIGMOID
    got = oracle.predict_sparse(imp.nodes, imp.tree_offsets, X, -np.inf, output=imp.output, global_bias=imp.global_bias)
    assert np.allclose(got, m.predict_proba(X)[:, 1], rtol=0, atol=2e-6)


def test_threshold_conversion_is_exact():
    thr = np.array([0.1, -0.1, 0.5, 1e-30, -3.0000001, 2.0 ** -130, 0.0], dtype=np.float64)
    t = importers._f32_strictly_above(thr)
    for a, b in zip(thr, t):
        below = np.nextafter(b, np.float32(-np.inf))
        assert float(below) <= a < float(b)          # x <= a  <=>  x < b  for every float32 x


def _xgb_fixture():
    # two trees in the `Booster.save_model("m.json")` layout; node 0 is the root, -1 = no child
    t0 = dict(left_children=[1, 3, -1, -1, -1], right_children=[2, 4, -1, -1, -1], split_indices=[2, 0, 0, 0, 0],
              split_conditions=[0.25, -0.5, 0.7, -0.3, 0.1], default_left=[1, 0, 0, 0, 0], sum_hessian=[10, 6, 4, 3, 3])
    t1 = dict(left_children=[1, -1, 3, -1, -1], right_children=[2, -1, 4, -1, -1], split_indices=[1, 0, 3, 0, 0],
              split_conditions=[0.0, 0.2, 0.5, -0.4, 0.9], default_left=[0, 0, 1, 0, 0], sum_hessian=[10, 5, 5, 2, 3])
    return {"learner": {"learner_model_param": {"base_score": "2.5E-1", "num_class": "0", "num_feature": "4"},
                        "objective": {"name": "binary:logistic"},
                        "gradient_booster": {"name": "gbtree", "model": {"trees": [t0, t1]}}}}


def _xgb_eval(model, x, missing):
    """XGBoost's documented rule, restated for the test: x < cond -> left, missing -> default_left."""
    s = 0.0
    for t in model["learner"]["gradient_booster"]["model"]["trees"]:
        n = 0
        while t["left_children"][n] != -1:
            v = x[t["split_indices"][n]]
            if abs(v - missing) <= 1e-6:
                left = bool(t["default_left"][n])
            else:
                left = bool(np.float32(v) < np.float32(t["split_conditions"][n]))
            n = t["left_children"][n] if left else t["right_children"][n]
        s += np.float32(t["split_conditions"][n])
    b = float(model["learner"]["learner_model_param"]["base_score"])
    return 1.0 / (1.0 + math.exp(-(s + math.log(b / (1 - b)))))


def test_xgboost_json(oracle, tmp_path):
    model = _xgb_fixture()
    path = tmp_path / "m.json"
    path.write_text(json.dumps(model))
    imp = importers.from_xgboost_json(str(path))
    assert (imp.num_trees, imp.num_cols, imp.output) == (2, 4, importers.OUT_SIGMOID)
    rng = np.random.default_rng(3)
    X = rng.uniform(-1, 1, size=(200, 4)).astype(np.float32)
    X[::7, 2] = -999.0; X[::5, 1] = -999.0; X[3, 2] = 0.25; X[4, 0] = -0.5; X[6, 3] = 0.5    # missing + ties
    got = oracle.predict_sparse(imp.nodes, imp.tree_offsets, X, -999.0, output=imp.output, global_bias=imp.global_bias)
    want = np.array([_xgb_eval(model, x, -999.0) for x in X])
    assert np.allclose(got, want, rtol=0, atol=1e-6)
    with pytest.raises(ValueError):
        bad = _xgb_fixture(); bad["learner"]["learner_model_param"]["num_class"] = "3"
        importers.from_xgboost_json(bad)


def test_imported_forest_packs_through_the_c_abi():
    X, y = _data(rows=300)
    m = sk.RandomForestRegressor(n_estimators=8, max_depth=5, random_state=1).fit(X, y)
    f = importers.from_sklearn(m).forest()
    inf = f.info()
    assert inf.num_trees == 8 and inf.num_nodes == sum(e.tree_.node_count for e in m.estimators_)


@pytest.mark.gpu
def test_sklearn_forest_on_the_gpu(oracle):
    import torch
    from tahoe_b200 import api
    X, y = _data(rows=5000, cols=20, seed=11)
    m = sk.RandomForestRegressor(n_estimators=64, max_depth=10, random_state=7).fit(X[:1500], y[:1500])
    Xt = _with_ties(list(m.estimators_), X)
    imp = importers.from_sklearn(m)
    f = imp.forest()
    want_leaves = np.stack([tr.apply(Xt) for tr in m.estimators_], axis=1)
    want = m.predict(Xt)
    x = torch.from_numpy(Xt).cuda()
    ran = 0
    for s in api.ALL_STRATEGIES:
        if not f.applicable(s, Xt.shape[0]):
            continue
        ids = torch.zeros((Xt.shape[0], imp.num_trees), dtype=torch.int32, device="cuda")
        got = f.predict(x, strategy=s, leaf_ids=ids).cpu().numpy()
        torch.cuda.synchronize()
        lids = ids.cpu().numpy()
        assert np.array_equal(_leaf_ids_as_source(imp, lids), want_leaves), api.STRATEGY_NAMES[s]
        assert np.all(np.abs(got - want) <= 1e-5 * np.maximum(_l1(imp, lids) / imp.num_trees, 1.0)), api.STRATEGY_NAMES[s]
        ran += 1
    assert ran >= 2


def _class_data(rows=900, cols=10, classes=4, seed=21):
    rng = np.random.default_rng(seed)
    X = rng.uniform(-1, 1, size=(rows, cols)).astype(np.float32)
    y = (np.floor((X[:, 0] + X[:, 1] * X[:, 2] + 1.5) / 3.0 * classes).clip(0, classes - 1)).astype(int)
    return X, y


def _oracle_classes(oracle, imps, X, missing=-np.inf):
    cols = [oracle.predict_sparse(i.nodes, i.tree_offsets, X, missing, output=i.output, global_bias=i.global_bias)
            for i in imps.per_class]
    out = np.stack(cols, axis=1).astype(np.float64)
    if imps.link == "softmax":
        out = np.exp(out - out.max(axis=1, keepdims=True))
        out /= out.sum(axis=1, keepdims=True)
    return out


@pytest.mark.parametrize("kind", ["rf", "gb"])
def test_sklearn_multiclass_per_class_outputs(oracle, kind):
    X, y = _class_data()
    if kind == "rf":
        m = sk.RandomForestClassifier(n_estimators=20, max_depth=6, random_state=3).fit(X, y)
    else:
        m = sk.GradientBoostingClassifier(n_estimators=15, max_depth=3, random_state=3).fit(X, y)
    imps = importers.from_sklearn_classes(m)
    assert imps.num_classes == 4
    got = _oracle_classes(oracle, imps, X)
    assert np.allclose(got, m.predict_proba(X), rtol=0, atol=5e-6)
    assert np.array_equal(got.argmax(1), m.predict(X))


def test_xgboost_multiclass_trees_go_to_their_class(oracle):
    base = _xgb_fixture()
    t0, t1 = base["learner"]["gradient_booster"]["model"]["trees"]
    model = {"learner": {"learner_model_param": {"base_score": "5E-1", "num_class": "3", "num_feature": "4"},
                         "objective": {"name": "multi:softprob"},
                         "gradient_booster": {"name": "gbtree", "model": {"trees": [t0, t1, t0, t1, t1, t0],
                                                                       "tree_info": [0, 1, 2, 0, 1, 2]}}}}
    imps = importers.from_xgboost_json_classes(model)
    assert [i.num_trees for i in imps.per_class] == [2, 2, 2] and imps.link == "softmax"
    # class 0 = trees {0, 3} = {t0, t1}: the same two trees as the binary fixture
    two = importers.from_xgboost_json(base)
    assert np.array_equal(imps.per_class[0].nodes, two.nodes)
    with pytest.raises(ValueError):
        importers.from_xgboost_json_classes(base)


@pytest.mark.gpu
def test_multiclass_forest_on_the_gpu(oracle):
    import torch
    X, y = _class_data(rows=4000, cols=16, classes=3, seed=5)
    m = sk.RandomForestClassifier(n_estimators=32, max_depth=8, random_state=1).fit(X[:1500], y[:1500])
    cf = importers.from_sklearn_classes(m).forests()
    got = cf.predict(torch.from_numpy(X).cuda()).cpu().numpy()
    assert got.shape == (4000, 3)
    assert np.allclose(got, m.predict_proba(X), rtol=0, atol=5e-6)
