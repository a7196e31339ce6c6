"""Importers: trained tree ensembles -> Tahoe compact nodes.

The reference loads forests only from its own text format, written by an external trainer repository
(README.md:32-33).  These importers go straight from a trained model to the compact (sparse) node
array `tahoe_forest_create_sparse` takes -- {val, bits, left_idx}, children adjacent
(Struct.h:50-54, 2244) -- with the reference's exact branch semantics (BaseTahoeTest.h:450-453):

    missing (|v - sentinel| <= 1e-6)  ->  default_left ? left : right
    otherwise                          ->  v >= val ? right : left           (NaN goes left)

* scikit-learn (`from_sklearn`): its trees test `x <= threshold -> left` on float32 features against
  a float64 threshold.  For float32 x that is exactly `x < t'` with t' = the smallest float32 above
  the threshold, so the converted forest selects the same leaves as `model.apply(X)`, row by row.
* XGBoost JSON (`from_xgboost_json`): `x < split_condition -> left`, missing -> `default_left`, all in
  float32 -- the reference's rule as it stands.  Leaf values are the `split_conditions` of leaf nodes;
  `dart` boosters scale every tree by its `weight_drop`; categorical splits are rejected.
* LightGBM text (`from_lightgbm_text`, the `Booster.save_model("m.txt")` format): `x <= threshold -> left`
  on double thresholds (same float32 conversion as scikit-learn); decision_type bit 1 = default_left,
  bits 2-3 = missing type (NaN: the default direction; None: LightGBM turns NaN into 0.0, so the node's
  default direction is where 0.0 goes; Zero: rejected -- "0.0 is missing" has no sentinel form).
* Treelite JSON (`from_treelite_json`, `Model.dump_as_json()`): numerical test nodes with
  `comparison_op` < / <= / > / >= (the last two by swapping the children), `default_left`, leaf values.

MISSING VALUES.  The trainers' missing value is NaN; the reference's is a finite sentinel from the data
header.  `Imported.forest()` therefore packs with the sentinel NaN by default, which this library reads
as "NaN is the missing value" (csrc/device_utils.cuh is_missing; the reference would match nothing).
Callers whose data marks absent features with a finite sentinel pass it as `missing=`; NaN inputs then
take the reference's path (every test fails: left).

Every importer returns an `Imported` record; `Imported.forest()` packs it through the C ABI.
"""
from __future__ import annotations

import json
import math
from dataclasses import dataclass, field

import numpy as np

from .synth import SPARSE_NODE

DEF_LEFT = 1 << 30           # Struct.h:58
IS_LEAF = 1 << 31            # Struct.h:59
OUT_RAW, OUT_AVG, OUT_SIGMOID, OUT_THRESHOLD = 0x0, 0x1, 0x10, 0x100   # Struct.h:37-42


@dataclass
class Imported:
    nodes: np.ndarray            # SPARSE_NODE [total]
    tree_offsets: np.ndarray     # int64 [T+1]
    weights: np.ndarray          # float32 [total]: training cover of every node (split-probability reorder)
    num_cols: int
    output: int = OUT_RAW
    global_bias: float = 0.0
    threshold: float = 0.0
    # per tree: compact index -> the trainer's own node id (compare leaf ids with e.g. model.apply)
    source_ids: list = field(default_factory=list)

    @property
    def num_trees(self) -> int:
        return int(self.tree_offsets.shape[0] - 1)

    def forest(self, missing: float = float("nan")):
        """Pack through the C ABI.  `missing` is the value the data uses for absent features: NaN (the
        default: the trainers' own convention -- NaN inputs take each node's default direction) or a
        finite sentinel (data-file header line 3 in the reference, BaseTahoeTest.h:370-371)."""
        if math.isinf(missing):
            raise ValueError("an infinite sentinel matches nothing (inf - inf is NaN): pass NaN or a finite value")
        from . import api
        return api.Forest.from_sparse(self.nodes, self.tree_offsets, self.num_cols, weights=self.weights,
                                      missing=missing, output=self.output, threshold=self.threshold,
                                      global_bias=self.global_bias)


def _f32_strictly_above(thr: np.ndarray) -> np.ndarray:
    """Smallest float32 t' with (x <= thr) == (x < t') for every float32 x."""
    thr = np.asarray(thr, dtype=np.float64)
    t32 = thr.astype(np.float32)
    over = t32.astype(np.float64) > thr                    # rounded up: step back below the threshold
    t32 = np.where(over, np.nextafter(t32, np.float32(-np.inf)), t32)
    return np.nextafter(t32, np.float32(np.inf)).astype(np.float32)


def _compact_tree(left, right, feature, split_val, leaf_val, def_left, cover):
    """One tree given as child-index arrays (-1 = leaf) -> compact nodes in breadth-first order with
    the two children of a node adjacent.  Returns (nodes, weights, source ids)."""
    order, left_idx, head = [0], [], 0
    while head < len(order):
        k = order[head]
        if left[k] < 0:
            left_idx.append(0)
        else:
            left_idx.append(len(order))
            order.append(int(left[k]))
            order.append(int(right[k]))
        head += 1
    order = np.asarray(order, dtype=np.int64)
    is_leaf = np.asarray(left)[order] < 0
    sp = np.zeros(len(order), dtype=SPARSE_NODE)
    sp["val"] = np.where(is_leaf, np.asarray(leaf_val, dtype=np.float32)[order],
                         np.asarray(split_val, dtype=np.float32)[order])
    fid = np.where(is_leaf, 0, np.asarray(feature)[order]).astype(np.int64)
    bits = fid | np.where(np.asarray(def_left, dtype=bool)[order] & ~is_leaf, DEF_LEFT, 0) | np.where(is_leaf, IS_LEAF, 0)
    sp["bits"] = (bits & 0xFFFFFFFF).astype(np.uint32).view(np.int32)
    sp["left_idx"] = np.asarray(left_idx, dtype=np.int32)
    return sp, np.asarray(cover, dtype=np.float32)[order], order.astype(np.int32)


def _assemble(trees, num_cols, **kw) -> Imported:
    nodes = np.concatenate([t[0] for t in trees])
    weights = np.concatenate([t[1] for t in trees])
    offs = np.zeros(len(trees) + 1, dtype=np.int64)
    offs[1:] = np.cumsum([len(t[0]) for t in trees])
    return Imported(nodes, offs, weights, int(num_cols), source_ids=[t[2] for t in trees], **kw)


# ---- scikit-learn ------------------------------------------------------------------------------
def from_sklearn(model) -> Imported:
    """RandomForestRegressor / ExtraTreesRegressor (mean of the trees: output AVG), DecisionTreeRegressor,
    GradientBoostingRegressor (init + learning_rate * sum: RAW + bias), GradientBoostingClassifier with
    two classes (the same raw score through a sigmoid = predict_proba[:, 1])."""
    name = type(model).__name__
    scale, bias, output = 1.0, 0.0, OUT_RAW
    if name in ("RandomForestRegressor", "ExtraTreesRegressor"):
        estimators, output = list(model.estimators_), OUT_AVG
    elif name == "DecisionTreeRegressor":
        estimators = [model]
    elif name in ("GradientBoostingRegressor", "GradientBoostingClassifier"):
        if model.estimators_.shape[1] != 1:
            raise ValueError("multi-class gradient boosting sums one forest per class; Tahoe's predict has one output")
        estimators = [e[0] for e in model.estimators_]
        scale = float(model.learning_rate)
        init = model.init_
        if init == "zero":
            bias = 0.0
        elif name == "GradientBoostingRegressor":
            bias = float(np.ravel(init.constant_)[0])
        else:                                                # log-odds of the class prior
            p1 = float(init.class_prior_[1])
            bias = math.log(p1 / (1.0 - p1))
            output = OUT_SIGMOID
    else:
        raise TypeError(f"from_sklearn: unsupported model {name}")
    trees = []
    for est in estimators:
        t = est.tree_
        if t.value.shape[1] != 1 or t.value.shape[2] != 1:
            raise ValueError("only single-output regression trees carry one value per leaf")
        left, right = t.children_left, t.children_right
        go_left = getattr(t, "missing_go_to_left", None)
        def_left = np.zeros(t.node_count, dtype=bool) if go_left is None else np.asarray(go_left, dtype=bool)
        split = _f32_strictly_above(np.where(left < 0, 0.0, t.threshold))
        trees.append(_compact_tree(left, right, np.maximum(t.feature, 0), split, t.value[:, 0, 0] * scale,
                                   def_left, t.weighted_n_node_samples))
    return _assemble(trees, model.n_features_in_, output=output, global_bias=bias)


# ---- XGBoost JSON --------------------------------------------------------------------------------
def from_xgboost_json(model) -> Imported:
    """`Booster.save_model("m.json")` schema: learner.gradient_booster.model.trees[*] with the arrays
    left_children / right_children (-1 = leaf), split_indices, split_conditions (threshold, or the leaf
    value of a leaf), default_left, sum_hessian.  Objectives: reg:* (RAW + base_score) and
    binary:logistic (sigmoid of the margin; base_score is a probability).  `model`: path or parsed dict."""
    if not isinstance(model, dict):
        with open(model) as f:
            model = json.load(f)
    learner = model["learner"]
    gb = learner["gradient_booster"]
    if gb.get("name", "gbtree") not in ("gbtree", "dart"):
        raise ValueError(f"unsupported booster {gb.get('name')}")
    gbm = gb["gbtree"]["model"] if "gbtree" in gb else gb["model"]
    lmp = learner["learner_model_param"]
    if int(lmp.get("num_class", "0")) > 1:
        raise ValueError("multi-class boosters sum one forest per class; Tahoe's predict has one output")
    objective = learner.get("objective", {}).get("name", "reg:squarederror")
    base_score = float(lmp.get("base_score", "0.5"))
    if objective == "binary:logistic":
        output, bias = OUT_SIGMOID, math.log(base_score / (1.0 - base_score))
    elif objective.startswith("reg:") and objective not in ("reg:gamma", "reg:tweedie", "reg:logistic"):
        output, bias = OUT_RAW, base_score
    else:
        raise ValueError(f"unsupported objective {objective}")
    drop = _xgb_weight_drop(gb, len(gbm["trees"]))
    trees = [_xgb_tree(t, drop[i]) for i, t in enumerate(gbm["trees"])]
    return _assemble(trees, int(lmp["num_feature"]), output=output, global_bias=bias)


def _xgb_weight_drop(gb, num_trees):
    """dart: prediction = sum_i weight_drop[i] * tree_i (gbtree: all ones)."""
    if gb.get("name", "gbtree") != "dart":
        return np.ones(num_trees, dtype=np.float64)
    w = gb.get("weight_drop")
    if w is None or len(w) != num_trees:
        raise ValueError("dart booster without a weight_drop entry per tree")
    return np.asarray(w, dtype=np.float64)


def _xgb_tree(t, scale=1.0):
    left = np.asarray(t["left_children"], dtype=np.int64)
    split_type = np.asarray(t.get("split_type", np.zeros(len(left))), dtype=np.int64)
    if np.any(split_type[left >= 0] != 0) or len(t.get("categories", [])) > 0:
        raise ValueError("categorical splits (split_type != 0) have no threshold form: not supported")
    cond = np.asarray(t["split_conditions"], dtype=np.float32)
    leaf = (cond.astype(np.float64) * scale).astype(np.float32) if scale != 1.0 else cond
    return _compact_tree(left, np.asarray(t["right_children"], dtype=np.int64),
                         np.asarray(t["split_indices"], dtype=np.int64), cond, leaf,
                         np.asarray(t["default_left"], dtype=bool),
                         np.asarray(t.get("sum_hessian", np.ones(len(left))), dtype=np.float32))


# ---- per-class outputs ---------------------------------------------
# The reference's predict has ONE output: it sums every tree into one scalar, also for multi-class
# models (BaseTahoeTest.h:462-466).  A K-class ensemble is K forests over the same rows -- one per
# class -- so it is composed above the C ABI: K handles, K predicts into the rows of a [K, R] buffer.
@dataclass
class ImportedClasses:
    per_class: list              # K Imported records, class k's trees in training order
    link: str = "raw"            # "raw": the K scores as they are; "softmax": over the K margins

    @property
    def num_classes(self) -> int:
        return len(self.per_class)

    def forests(self, missing: float = float("nan")):
        """K handles, one per class (K quantise passes per batch): kept as the cross-check of `forest()`."""
        return ClassForests([imp.forest(missing) for imp in self.per_class], self.link)

    def forest(self, missing: float = float("nan")):
        """ONE packed forest for the K classes through the C ABI (tahoe_forest_create_sparse_classes): one set of
        rank tables, so one quantise pass per batch serves every class.  The classes must share their output
        transform; a per-class bias (multi-class gradient boosting priors) is added to the K scores afterwards."""
        from . import api
        if math.isinf(missing):
            raise ValueError("an infinite sentinel matches nothing (inf - inf is NaN): pass NaN or a finite value")
        first = self.per_class[0]
        if any(i.output != first.output or i.threshold != first.threshold or i.num_cols != first.num_cols
               for i in self.per_class):
            raise ValueError("the classes of one packed forest share output transform and width")
        if (first.output & (OUT_SIGMOID | OUT_THRESHOLD)) and any(i.global_bias != first.global_bias for i in self.per_class):
            raise ValueError("per-class biases do not commute with a sigmoid / threshold transform")
        nodes = np.concatenate([i.nodes for i in self.per_class])
        weights = np.concatenate([i.weights for i in self.per_class])
        sizes = np.concatenate([np.diff(i.tree_offsets) for i in self.per_class])
        offs = np.zeros(sizes.shape[0] + 1, dtype=np.int64)
        offs[1:] = np.cumsum(sizes)
        tree_class = np.concatenate([np.full(i.num_trees, k, dtype=np.int32) for k, i in enumerate(self.per_class)])
        common_bias = first.global_bias if all(i.global_bias == first.global_bias for i in self.per_class) else 0.0
        f = api.Forest.from_sparse_classes(nodes, offs, first.num_cols, tree_class, self.num_classes, weights=weights,
                                           missing=missing, output=first.output, threshold=first.threshold,
                                           global_bias=common_bias)
        extra = None if common_bias == first.global_bias and all(i.global_bias == common_bias for i in self.per_class) \
            else np.array([i.global_bias for i in self.per_class], dtype=np.float32)
        return PackedClassForest(f, self.link, extra)


class ClassForests:
    """K packed forests predicted into one [R, K] tensor (device rows in, device scores out)."""

    def __init__(self, forests, link="raw"):
        self.forests, self.link = forests, link

    def predict(self, data, *, strategy=None):
        import torch
        from . import api
        rows = data.shape[0]
        out = torch.empty((len(self.forests), rows), dtype=torch.float32, device=data.device)
        for k, f in enumerate(self.forests):
            f.predict(data, out[k], strategy=api.AUTO if strategy is None else strategy)
        out = out.t()
        return torch.softmax(out, dim=1) if self.link == "softmax" else out


class PackedClassForest:
    """One handle for K classes: predict -> [R, K] (device rows in, device scores out)."""

    def __init__(self, forest, link="raw", class_bias=None):
        self.forest, self.link, self.class_bias = forest, link, class_bias

    def predict(self, data, *, strategy=None):
        import torch
        from . import api
        out = self.forest.predict_classes(data, strategy=api.AUTO if strategy is None else strategy).t()
        if self.class_bias is not None:
            out = out + torch.from_numpy(self.class_bias).to(out.device)[None, :]
        return torch.softmax(out, dim=1) if self.link == "softmax" else out


def from_sklearn_classes(model) -> ImportedClasses:
    """RandomForestClassifier / ExtraTreesClassifier (class k's score = mean over trees of the leaf's
    class-k fraction = predict_proba[:, k]) and multi-class GradientBoostingClassifier (one regression
    tree per class and stage; softmax over init + learning_rate * sums = predict_proba)."""
    name = type(model).__name__
    K = int(model.n_classes_)
    per_class = []
    if name in ("RandomForestClassifier", "ExtraTreesClassifier"):
        for k in range(K):
            trees = []
            for est in model.estimators_:
                t = est.tree_
                left = t.children_left
                go_left = getattr(t, "missing_go_to_left", None)
                def_left = np.zeros(t.node_count, dtype=bool) if go_left is None else np.asarray(go_left, dtype=bool)
                val = t.value[:, 0, :]
                frac = val[:, k] / np.maximum(val.sum(axis=1), 1e-300)      # counts or fractions: normalise
                split = _f32_strictly_above(np.where(left < 0, 0.0, t.threshold))
                trees.append(_compact_tree(left, t.children_right, np.maximum(t.feature, 0), split, frac,
                                           def_left, t.weighted_n_node_samples))
            per_class.append(_assemble(trees, model.n_features_in_, output=OUT_AVG))
        return ImportedClasses(per_class, "raw")
    if name == "GradientBoostingClassifier" and model.estimators_.shape[1] == K and K > 2:
        prior = np.log(np.asarray(model.init_.class_prior_, dtype=np.float64)) if model.init_ != "zero" else np.zeros(K)
        for k in range(K):
            trees = []
            for est in model.estimators_[:, k]:
                t = est.tree_
                left = t.children_left
                go_left = getattr(t, "missing_go_to_left", None)
                def_left = np.zeros(t.node_count, dtype=bool) if go_left is None else np.asarray(go_left, dtype=bool)
                split = _f32_strictly_above(np.where(left < 0, 0.0, t.threshold))
                trees.append(_compact_tree(left, t.children_right, np.maximum(t.feature, 0), split,
                                           t.value[:, 0, 0] * float(model.learning_rate), def_left,
                                           t.weighted_n_node_samples))
            per_class.append(_assemble(trees, model.n_features_in_, output=OUT_RAW, global_bias=float(prior[k])))
        return ImportedClasses(per_class, "softmax")
    raise TypeError(f"from_sklearn_classes: unsupported model {name}")


def from_xgboost_json_classes(model) -> ImportedClasses:
    """multi:softprob / multi:softmax boosters: tree i belongs to class `tree_info[i]`; the K margins
    (base_score + sums) go through a softmax."""
    if not isinstance(model, dict):
        with open(model) as f:
            model = json.load(f)
    learner = model["learner"]
    gb = learner["gradient_booster"]
    gbm = gb["gbtree"]["model"] if "gbtree" in gb else gb["model"]
    lmp = learner["learner_model_param"]
    K = int(lmp.get("num_class", "0"))
    if K < 2:
        raise ValueError("not a multi-class booster: use from_xgboost_json")
    base_score = float(lmp.get("base_score", "0.5"))
    info = gbm.get("tree_info") or [i % K for i in range(len(gbm["trees"]))]
    drop = _xgb_weight_drop(gb, len(gbm["trees"]))
    per_class = []
    for k in range(K):
        trees = [_xgb_tree(t, drop[i]) for i, (t, cls) in enumerate(zip(gbm["trees"], info)) if int(cls) == k]
        per_class.append(_assemble(trees, int(lmp["num_feature"]), output=OUT_RAW, global_bias=base_score))
    return ImportedClasses(per_class, "softmax")


# ---- LightGBM model text ---------------------------------------------------------------------------
def _lgb_parse(text: str):
    """-> (header dict, list of per-tree dicts of raw strings)."""
    header, trees, cur = {}, [], None
    for line in text.splitlines():
        line = line.strip()
        if not line:
            continue
        if line == "end of trees":
            break
        if line.startswith("Tree="):
            cur = {}
            trees.append(cur)
            continue
        if "=" in line:
            k, v = line.split("=", 1)
            (header if cur is None else cur)[k] = v
        elif cur is None:
            header[line] = ""          # flags such as `average_output`
    return header, trees


def _lgb_tree(t):
    """One LightGBM tree -> compact nodes.  Internal nodes 0..n-2, leaves addressed as ~index
    (child < 0 -> leaf -child-1); rule `x <= threshold -> left`; decision_type: bit 0 categorical,
    bit 1 default_left, bits 2-3 missing type (0 none, 1 zero, 2 NaN)."""
    nl = int(t["num_leaves"])
    leaf_value = np.asarray(t["leaf_value"].split(), dtype=np.float64)   # already carries the shrinkage
    leaf_cnt = np.asarray(t.get("leaf_count", " ".join(["1"] * nl)).split(), dtype=np.float64)
    if nl == 1:                                         # a stump: one leaf, no split arrays
        return _compact_tree(np.array([-1]), np.array([-1]), np.array([0]), np.zeros(1, np.float32),
                             leaf_value[:1], np.zeros(1, bool), leaf_cnt[:1])
    ni = nl - 1
    feat = np.asarray(t["split_feature"].split(), dtype=np.int64)
    thr = np.asarray(t["threshold"].split(), dtype=np.float64)
    dec = np.asarray(t["decision_type"].split(), dtype=np.int64)
    lc = np.asarray(t["left_child"].split(), dtype=np.int64)
    rc = np.asarray(t["right_child"].split(), dtype=np.int64)
    int_cnt = np.asarray(t.get("internal_count", " ".join(["1"] * ni)).split(), dtype=np.float64)
    if np.any(dec & 1):
        raise ValueError("categorical splits (decision_type bit 0) have no threshold form: not supported")
    mtype = (dec >> 2) & 3
    if np.any(mtype == 1):
        raise ValueError("missing type Zero (0.0 is a missing value) has no sentinel form: retrain with "
                         "zero_as_missing=false")
    # NaN: the stored default direction; None: LightGBM turns NaN into 0.0, which goes left iff 0.0 <= threshold
    def_left = np.where(mtype == 2, (dec & 2) != 0, 0.0 <= thr)
    unify = lambda c: np.where(c >= 0, c, ni + (-c - 1))       # noqa: E731  leaves after the internal nodes
    left = np.concatenate([unify(lc), -np.ones(nl, np.int64)])
    right = np.concatenate([unify(rc), -np.ones(nl, np.int64)])
    z = np.zeros(nl)
    return _compact_tree(left, right, np.concatenate([feat, z.astype(np.int64)]),
                         np.concatenate([_f32_strictly_above(thr), z.astype(np.float32)]),
                         np.concatenate([np.zeros(ni), leaf_value]),
                         np.concatenate([def_left, z.astype(bool)]), np.concatenate([int_cnt, leaf_cnt]))


def _lgb_load(model):
    text = model
    if "\n" not in model:
        with open(model) as f:
            text = f.read()
    header, trees = _lgb_parse(text)
    if not trees:
        raise ValueError("no `Tree=` sections: not a LightGBM model text")
    return header, trees


def from_lightgbm_text(model) -> Imported:
    """LightGBM `Booster.save_model()` text (a path, or the text itself).  Objectives: regression*
    (raw sum), binary with sigmoid:1 (sigmoid of the sum); `average_output` (random-forest mode) -> AVG."""
    header, trees = _lgb_load(model)
    if int(header.get("num_tree_per_iteration", "1")) != 1:
        raise ValueError("multi-class model: use from_lightgbm_text_classes")
    obj = header.get("objective", "regression").split()
    output = OUT_RAW
    if obj[0] == "binary":
        sig = [o for o in obj[1:] if o.startswith("sigmoid:")]
        if sig and float(sig[0].split(":")[1]) != 1.0:
            raise ValueError("binary objective with sigmoid != 1 is not a plain logistic link")
        output = OUT_SIGMOID
    elif not (obj[0].startswith("regression") or obj[0] in ("huber", "fair", "quantile", "mape", "custom")):
        raise ValueError(f"unsupported objective {obj[0]}")
    if "average_output" in header:
        output |= OUT_AVG
    return _assemble([_lgb_tree(t) for t in trees], int(header["max_feature_idx"]) + 1, output=output)


def from_lightgbm_text_classes(model) -> ImportedClasses:
    """multiclass / multiclassova LightGBM models: tree i belongs to class i % num_tree_per_iteration."""
    header, trees = _lgb_load(model)
    K = int(header.get("num_tree_per_iteration", "1"))
    if K < 2:
        raise ValueError("not a multi-class model: use from_lightgbm_text")
    obj = header.get("objective", "multiclass").split()[0]
    link = "softmax" if obj == "multiclass" else "raw"
    cols = int(header["max_feature_idx"]) + 1
    out = OUT_SIGMOID if obj == "multiclassova" else OUT_RAW
    return ImportedClasses([_assemble([_lgb_tree(t) for t in trees[k::K]], cols, output=out) for k in range(K)], link)


# ---- Treelite JSON ---------------------------------------------------------------------------------
def from_treelite_json(model) -> Imported:
    """`treelite.Model.dump_as_json()` (a path, the JSON text, or the parsed dict): trees of
    numerical_test_node {split_feature_id, default_left, comparison_op, threshold, left_child, right_child}
    and leaf nodes {leaf_value}.  `x op threshold` true -> left child.  Single-output models only."""
    if not isinstance(model, dict):
        text = model
        if not model.lstrip().startswith("{"):
            with open(model) as f:
                text = f.read()
        model = json.loads(text)
    mp = model.get("model_param", {})
    post = model.get("postprocessor", mp.get("pred_transform", "identity"))
    if post in ("identity", "identity_multiclass"):
        output = OUT_RAW
    elif post == "sigmoid":
        if float(model.get("sigmoid_alpha", mp.get("sigmoid_alpha", 1.0))) != 1.0:
            raise ValueError("sigmoid_alpha != 1 is not a plain logistic link")
        output = OUT_SIGMOID
    else:
        raise ValueError(f"unsupported postprocessor {post}")
    if model.get("average_tree_output", False):
        output |= OUT_AVG
    nclass = model.get("num_class", [1])
    if (max(nclass) if isinstance(nclass, list) else int(nclass)) > 1 or int(model.get("num_target", 1)) > 1:
        raise ValueError("multi-class / multi-target Treelite models sum one forest per output; Tahoe's predict has one")
    base = model.get("base_scores", [mp.get("global_bias", 0.0)])
    trees = []
    for t in model["trees"]:
        nodes = {int(n["node_id"]): n for n in t["nodes"]}
        ids = sorted(nodes)
        root = int(t.get("root_id", ids[0]))
        if ids[0] != root:                             # _compact_tree starts at index 0
            ids.remove(root)
            ids.insert(0, root)
        pos = {nid: i for i, nid in enumerate(ids)}
        n = len(ids)
        left, right = -np.ones(n, np.int64), -np.ones(n, np.int64)
        feat, thr = np.zeros(n, np.int64), np.zeros(n, np.float32)
        leaf, dl, cover = np.zeros(n, np.float64), np.zeros(n, bool), np.ones(n, np.float64)
        for nid in ids:
            nd, i = nodes[nid], pos[nid]
            cover[i] = float(nd.get("data_count", nd.get("sum_hess", 1.0)))
            if "leaf_value" in nd:
                v = nd["leaf_value"]
                if isinstance(v, list):
                    if len(v) != 1:
                        raise ValueError("leaf vectors: one forest per output is needed")
                    v = v[0]
                leaf[i] = float(v)
                continue
            if nd.get("node_type", "numerical_test_node") != "numerical_test_node":
                raise ValueError("categorical test nodes have no threshold form: not supported")
            op = nd["comparison_op"]
            lch, rch, d = pos[int(nd["left_child"])], pos[int(nd["right_child"])], bool(nd["default_left"])
            if op in (">", ">="):                      # x > t -> left   ==   x <= t -> (old) right: swap the children
                lch, rch, d = rch, lch, not d
                op = "<=" if op == ">" else "<"
            t64 = float(nd["threshold"])
            if op == "<":                              # x < t  <=>  x < first float32 >= t
                t32 = np.float32(t64)
                thr[i] = t32 if float(t32) >= t64 else np.nextafter(t32, np.float32(np.inf))
            elif op == "<=":
                thr[i] = _f32_strictly_above(np.array([t64]))[0]
            else:
                raise ValueError(f"unsupported comparison_op {op}")
            left[i], right[i], feat[i], dl[i] = lch, rch, int(nd["split_feature_id"]), d
        sp, w, order = _compact_tree(left, right, feat, thr, leaf, dl, cover)
        trees.append((sp, w, np.asarray(ids, dtype=np.int32)[order]))      # source ids = node_id
    return _assemble(trees, int(model["num_feature"]), output=output, global_bias=float(base[0]))
