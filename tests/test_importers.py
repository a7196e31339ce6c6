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


def _oracle_classes(oracle, imps, X, missing=-np.inf):   # (the oracle takes any sentinel; -inf matches nothing)
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


# ---- LightGBM model text / Treelite JSON / dart / NaN as the missing value ---------------------------
# lightgbm, treelite and xgboost are not installed in this image, so these formats are pinned by
# hand-written fixtures in the trainers' documented layouts and by evaluators that restate their documented
# decision rules; scikit-learn IS installed, and its own predict on NaN inputs pins the NaN path.
_LGB_TEXT = """tree
version=v4
num_class=1
num_tree_per_iteration=1
label_index=0
max_feature_idx=3
objective=binary sigmoid:1
feature_names=f0 f1 f2 f3
feature_infos=[-1:1] [-1:1] [-1:1] [-1:1]
tree_sizes=400 300 120

Tree=0
num_leaves=4
num_cat=0
split_feature=2 0 3
split_gain=10 5 2
threshold=0.25000000000000006 -0.5 0.10000000000000002
decision_type=2 8 10
left_child=1 -1 -3
right_child=2 -2 -4
leaf_value=0.11 -0.22 0.33 -0.44
leaf_weight=1 1 1 1
leaf_count=30 20 25 25
internal_value=0 0 0
internal_weight=0 0 0
internal_count=100 50 50
is_linear=0
shrinkage=1

Tree=1
num_leaves=3
num_cat=0
split_feature=1 1
split_gain=3 1
threshold=0 0.5
decision_type=0 10
left_child=-1 -2
right_child=1 -3
leaf_value=0.05 0.15 -0.25
leaf_weight=1 1 1
leaf_count=40 35 25
internal_value=0 0
internal_weight=0 0
internal_count=100 60
is_linear=0
shrinkage=0.1

Tree=2
num_leaves=1
num_cat=0
leaf_value=0.0625
is_linear=0
shrinkage=0.1

end of trees

feature_importances:
f1=2
"""


def _lgb_eval(text, x):
    """LightGBM's documented numerical decision, restated: missing type NaN -> NaN takes default_left;
    missing type None -> NaN is read as 0.0; then x <= threshold (double) -> left.  Leaves are ~index."""
    header, trees = importers._lgb_parse(text)
    s = 0.0
    for t in trees:
        lv = [float(v) for v in t["leaf_value"].split()]
        if int(t["num_leaves"]) == 1:
            s += lv[0]
            continue
        feat = [int(v) for v in t["split_feature"].split()]
        thr = [float(v) for v in t["threshold"].split()]
        dec = [int(v) for v in t["decision_type"].split()]
        lc = [int(v) for v in t["left_child"].split()]
        rc = [int(v) for v in t["right_child"].split()]
        n = 0
        while n >= 0:
            v = float(x[feat[n]])
            mt = (dec[n] >> 2) & 3
            if math.isnan(v) and mt == 2:
                left = bool(dec[n] & 2)
            else:
                if math.isnan(v):
                    v = 0.0
                left = v <= thr[n]
            n = lc[n] if left else rc[n]
        s += lv[-n - 1]
    return 1.0 / (1.0 + math.exp(-s))


def _nan_to_sentinel(X, sentinel=-999.0):
    Y = X.copy()
    Y[np.isnan(Y)] = sentinel
    return Y


def _nan_data(rows=400, cols=4, seed=13):
    rng = np.random.default_rng(seed)
    X = rng.uniform(-1, 1, size=(rows, cols)).astype(np.float32)
    X[rng.uniform(size=X.shape) < 0.15] = np.nan
    return X


def test_lightgbm_text(oracle, tmp_path):
    imp = importers.from_lightgbm_text(_LGB_TEXT)
    assert (imp.num_trees, imp.num_cols, imp.output) == (3, 4, importers.OUT_SIGMOID)
    assert [len(sid) for sid in imp.source_ids] == [7, 5, 1]
    X = _nan_data()
    X[3, 2] = np.float32(0.25); X[4, 0] = -0.5; X[5, 1] = 0.0; X[6, 1] = 0.5; X[7, 3] = np.float32(0.1)   # ties
    # the oracle knows finite sentinels only: NaN -> -999 with sentinel -999 is the same rule
    got = oracle.predict_sparse(imp.nodes, imp.tree_offsets, _nan_to_sentinel(X), -999.0, output=imp.output,
                                global_bias=imp.global_bias)
    want = np.array([_lgb_eval(_LGB_TEXT, x) for x in X])
    assert np.allclose(got, want, rtol=0, atol=1e-6)
    path = tmp_path / "m.txt"
    path.write_text(_LGB_TEXT)
    assert np.array_equal(importers.from_lightgbm_text(str(path)).nodes, imp.nodes)
    with pytest.raises(ValueError):      # zero_as_missing
        importers.from_lightgbm_text(_LGB_TEXT.replace("decision_type=2 8 10", "decision_type=2 4 10"))
    with pytest.raises(ValueError):      # categorical
        importers.from_lightgbm_text(_LGB_TEXT.replace("decision_type=2 8 10", "decision_type=3 8 10"))
    with pytest.raises(ValueError):
        importers.from_lightgbm_text_classes(_LGB_TEXT)
    multi = _LGB_TEXT.replace("num_tree_per_iteration=1", "num_tree_per_iteration=3").replace(
        "objective=binary sigmoid:1", "objective=multiclass num_class:3")
    mc = importers.from_lightgbm_text_classes(multi)
    assert mc.num_classes == 3 and mc.link == "softmax" and [i.num_trees for i in mc.per_class] == [1, 1, 1]


def _treelite_fixture():
    n = lambda i, f, d, op, t, l, r: dict(node_id=i, split_feature_id=f, default_left=d, node_type="numerical_test_node",  # noqa: E731
                                          comparison_op=op, threshold=t, left_child=l, right_child=r, data_count=10)
    leaf = lambda i, v: dict(node_id=i, leaf_value=v, data_count=5)     # noqa: E731
    t0 = dict(num_nodes=7, has_categorical_split=False, root_id=0,
              nodes=[n(0, 2, True, "<", 0.25, 1, 2), n(1, 0, False, "<=", -0.5, 3, 4), n(2, 3, True, ">", 0.1, 5, 6),
                     leaf(3, 0.5), leaf(4, -0.25), leaf(5, 0.125), leaf(6, -1.0)])
    t1 = dict(num_nodes=3, has_categorical_split=False, root_id=0,
              nodes=[n(0, 1, False, ">=", 0.0, 1, 2), leaf(1, [0.75]), leaf(2, -0.5)])
    return {"num_feature": 4, "task_type": "kRegressor", "average_tree_output": True, "num_target": 1, "num_class": [1],
            "leaf_vector_shape": [1, 1], "postprocessor": "identity", "sigmoid_alpha": 1.0, "base_scores": [0.5],
            "trees": [t0, t1]}


def _treelite_eval(model, x):
    ops = {"<": lambda a, b: a < b, "<=": lambda a, b: a <= b, ">": lambda a, b: a > b, ">=": lambda a, b: a >= b}
    s = 0.0
    for t in model["trees"]:
        nodes = {n["node_id"]: n for n in t["nodes"]}
        n = nodes[t["root_id"]]
        while "leaf_value" not in n:
            v = float(x[n["split_feature_id"]])
            left = n["default_left"] if math.isnan(v) else ops[n["comparison_op"]](v, n["threshold"])
            n = nodes[n["left_child"] if left else n["right_child"]]
        lv = n["leaf_value"]
        s += lv[0] if isinstance(lv, list) else lv
    return s / len(model["trees"]) + model["base_scores"][0]


def test_treelite_json(oracle):
    model = _treelite_fixture()
    imp = importers.from_treelite_json(json.dumps(model))
    assert (imp.num_trees, imp.num_cols, imp.output, imp.global_bias) == (2, 4, importers.OUT_AVG, 0.5)
    X = _nan_data(seed=17)
    X[3, 2] = 0.25; X[4, 0] = -0.5; X[5, 3] = np.float32(0.1); X[6, 1] = 0.0      # ties, every operator
    got, lids = oracle.predict_sparse(imp.nodes, imp.tree_offsets, _nan_to_sentinel(X), -999.0, output=imp.output,
                                      global_bias=imp.global_bias, want_leaf_ids=True)
    want = np.array([_treelite_eval(model, x) for x in X])
    assert np.allclose(got, want, rtol=0, atol=1e-6)
    assert set(np.unique(_leaf_ids_as_source(imp, lids)[:, 0])) <= {3, 4, 5, 6}   # leaf ids map back to node_id
    bad = _treelite_fixture(); bad["trees"][0]["nodes"][0]["node_type"] = "categorical_test_node"
    with pytest.raises(ValueError):
        importers.from_treelite_json(bad)
    bad = _treelite_fixture(); bad["num_class"] = [3]
    with pytest.raises(ValueError):
        importers.from_treelite_json(bad)


def test_xgboost_dart_and_categorical(oracle):
    model = _xgb_fixture()
    gb = model["learner"]["gradient_booster"]
    model["learner"]["gradient_booster"] = {"name": "dart", "gbtree": {"model": gb["model"]}, "weight_drop": [0.5, 0.25]}
    imp = importers.from_xgboost_json(model)
    plain = importers.from_xgboost_json(_xgb_fixture())
    leaf = plain.nodes["bits"] < 0
    scale = np.where(np.arange(plain.nodes.shape[0]) < plain.tree_offsets[1], 0.5, 0.25)
    assert np.array_equal(imp.nodes["val"][leaf], (plain.nodes["val"][leaf].astype(np.float64) * scale[leaf]).astype(np.float32))
    assert np.array_equal(imp.nodes["val"][~leaf], plain.nodes["val"][~leaf])       # thresholds untouched
    del model["learner"]["gradient_booster"]["weight_drop"]
    with pytest.raises(ValueError):
        importers.from_xgboost_json(model)
    cat = _xgb_fixture()
    cat["learner"]["gradient_booster"]["model"]["trees"][0]["split_type"] = [1, 0, 0, 0, 0]
    with pytest.raises(ValueError):
        importers.from_xgboost_json(cat)


def test_sklearn_nan_inputs_take_the_trained_default_direction(oracle):
    """scikit-learn trees trained on data with NaN carry missing_go_to_left; the imported forest packed with
    the NaN sentinel must route NaN inputs the same way (here through the oracle: NaN -> -999, sentinel -999)."""
    X, y = _data(rows=800, cols=8, seed=31)
    rng = np.random.default_rng(2)
    Xn = X.copy()
    Xn[rng.uniform(size=X.shape) < 0.1] = np.nan
    m = sk.RandomForestRegressor(n_estimators=12, max_depth=6, random_state=5).fit(Xn, y)
    imp = importers.from_sklearn(m)
    got, lids = oracle.predict_sparse(imp.nodes, imp.tree_offsets, _nan_to_sentinel(Xn), -999.0, output=imp.output,
                                      global_bias=imp.global_bias, want_leaf_ids=True)
    assert np.array_equal(_leaf_ids_as_source(imp, lids), np.stack([tr.apply(Xn) for tr in m.estimators_], axis=1))
    assert np.allclose(got, m.predict(Xn), rtol=0, atol=1e-5)
    with pytest.raises(ValueError):
        imp.forest(missing=float("-inf"))


@pytest.mark.gpu
def test_nan_is_the_missing_value_on_the_gpu(oracle):
    """Imported.forest() packs with the NaN sentinel: every strategy must route NaN features by default_left
    (csrc/device_utils.cuh is_missing) -- against scikit-learn's own apply/predict on NaN inputs and against the
    LightGBM / Treelite fixtures' evaluators."""
    import torch
    from tahoe_b200 import api
    X, y = _data(rows=6000, cols=12, seed=41)
    rng = np.random.default_rng(3)
    Xn = X.copy()
    Xn[rng.uniform(size=X.shape) < 0.1] = np.nan
    m = sk.RandomForestRegressor(n_estimators=48, max_depth=9, random_state=6).fit(Xn[:2000], y[:2000])
    imp = importers.from_sklearn(m)
    f = imp.forest()                                    # sentinel NaN
    want_leaves = np.stack([tr.apply(Xn) for tr in m.estimators_], axis=1)
    want = m.predict(Xn)
    x = torch.from_numpy(Xn).cuda()
    ran = 0
    for s in api.ALL_STRATEGIES:
        if not f.applicable(s, Xn.shape[0]):
            continue
        ids = torch.zeros((Xn.shape[0], imp.num_trees), dtype=torch.int32, device="cuda")
        got = f.predict(x, strategy=s, leaf_ids=ids).cpu().numpy()
        lids = ids.cpu().numpy()
        assert np.array_equal(_leaf_ids_as_source(imp, lids), want_leaves), api.STRATEGY_NAMES[s]
        tol = 1e-5 * np.maximum(_l1(imp, lids) / imp.num_trees, 1.0)
        assert np.all(np.abs(got - want) <= tol), api.STRATEGY_NAMES[s]
        got2 = f.predict(x, strategy=s).cpu().numpy()  # the performance instantiation
        assert np.all(np.abs(got2 - want) <= tol), api.STRATEGY_NAMES[s]
        ran += 1
    assert ran >= 2
    lgb = importers.from_lightgbm_text(_LGB_TEXT)
    Xl = _nan_data(rows=3000, seed=19)
    got = lgb.forest().predict(torch.from_numpy(Xl).cuda()).cpu().numpy()
    assert np.allclose(got, np.array([_lgb_eval(_LGB_TEXT, r) for r in Xl]), rtol=0, atol=1e-6)
    tl = importers.from_treelite_json(_treelite_fixture())
    got = tl.forest().predict(torch.from_numpy(Xl).cuda()).cpu().numpy()
    assert np.allclose(got, np.array([_treelite_eval(_treelite_fixture(), r) for r in Xl]), rtol=0, atol=1e-6)


def test_class_forest_packs_through_the_c_abi():
    """tahoe_forest_create_sparse_classes: classes padded to whole traversal groups with single-leaf trees."""
    X, y = _class_data(rows=600)
    m = sk.RandomForestClassifier(n_estimators=10, max_depth=5, random_state=2).fit(X, y)
    pf = importers.from_sklearn_classes(m).forest()
    inf = pf.forest.info()
    assert pf.forest.num_classes == 4 and inf.num_trees == 4 * 12            # 10 real + 2 padding trees per class
    assert inf.num_nodes == 4 * (sum(e.tree_.node_count for e in m.estimators_) + 2)


@pytest.mark.gpu
def test_per_class_outputs_in_one_handle_on_the_gpu(oracle):
    """tahoe_predict_classes (one packed forest, ONE quantise pass) against K separate handles and against the
    trainer's own predict_proba, through every strategy it supports, incl. NaN inputs."""
    import torch
    from tahoe_b200 import api
    X, y = _class_data(rows=20000, cols=16, classes=3, seed=5)
    rng = np.random.default_rng(8)
    X[rng.uniform(size=X.shape) < 0.05] = np.nan
    m = sk.RandomForestClassifier(n_estimators=50, max_depth=9, random_state=1).fit(X[:3000], y[:3000])
    imps = importers.from_sklearn_classes(m)
    packed, separate = imps.forest(), imps.forests()
    x = torch.from_numpy(X).cuda()
    want = m.predict_proba(X)
    sep = separate.predict(x, strategy=api.DIRECT).cpu().numpy()
    assert np.allclose(sep, want, rtol=0, atol=5e-6)
    ran = 0
    for s in (api.DIRECT, api.SHARED_DATA_PER_THREAD, api.SPLIT_SHARED_FOREST, api.AUTO):
        if s != api.AUTO and not packed.forest.applicable(s, X.shape[0]):
            continue
        got = packed.predict(x, strategy=s).cpu().numpy()
        assert got.shape == (X.shape[0], 3)
        assert np.allclose(got, want, rtol=0, atol=5e-6), s
        if s in (api.DIRECT, api.SHARED_DATA_PER_THREAD):                     # same trees in the same order: same bits
            assert np.array_equal(got.view(np.uint32), sep.view(np.uint32)), s
        if s == api.SPLIT_SHARED_FOREST:                                      # one k_quantize + one traverse per class
            assert packed.forest.launches() == 1 + 3
        ran += 1
    assert ran >= 3
    # multi-class gradient boosting: per-class priors, softmax link; a small batch (fewer rows than one wave)
    Xs, ys = _class_data(rows=700, cols=10, classes=4, seed=21)
    g = sk.GradientBoostingClassifier(n_estimators=12, max_depth=3, random_state=3).fit(Xs, ys)
    pg = importers.from_sklearn_classes(g).forest()
    for s in (api.DIRECT, api.SPLIT_SHARED_FOREST):
        got = pg.predict(torch.from_numpy(Xs).cuda(), strategy=s).cpu().numpy()
        assert np.allclose(got, g.predict_proba(Xs), rtol=0, atol=5e-6), s
