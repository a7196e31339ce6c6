"""Shared helpers for the parity tests."""
import os

import numpy as np

from tahoe_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_CASES = ["tiny", "char_w", "short_w", "int_w", "ties", "one_feature", "depth0",
                "susy_like", "higgs_like", "svhn_like"]
TRANSFORMS = [(0x0, 0.0, 0.0), (0x1, 0.0, 0.0), (0x11, 0.0, 0.25), (0x110, 0.5, -0.1),
              (0x111, 0.45, 0.0), (0x0, 0.0, 1.5)]


def load_case(name):
    """Golden case -> dict with nodes/data regenerated from seeds when not stored."""
    z = dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))
    c = {k: (v.item() if v.ndim == 0 else v) for k, v in z.items()}
    if "nodes" not in c:
        c["nodes"] = synth.gen_dense_forest(c["forest_seed"], c["trees"], c["depth"], c["cols"],
                                            leaf_prob=c["leaf_prob"], grid=c["grid"])
        c["data"] = synth.gen_data(c["data_seed"], c["rows"], c["cols"], miss_frac=c["miss_frac"],
                                   nan_frac=c["nan_frac"], missing=float(c["missing"]),
                                   grid=c["grid"])
    return c


def golden_preds(c, flags, thr, bias):
    return c[f"preds_{flags:03x}_{thr}_{bias}"]


def bits_equal(a, b):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


def l1_mass(nodes, num_trees, depth, leaf_ids):
    """sum_t |leaf value reached in tree t| per row, from the oracle's leaf ids (dense forests)."""
    npt = (1 << (depth + 1)) - 1
    vals = nodes["val"].reshape(num_trees, npt)
    t = np.arange(num_trees)[None, :]
    return np.abs(vals[t, leaf_ids].astype(np.float64)).sum(axis=1)


def score_close(got, want, l1=None):
    """north_star tolerance: summed scores within 1e-5 relative.  For a float sum whose order
    differs, "relative" is relative to the L1 mass of the summands: a sum of 500 leaf values that
    cancels to ~0 carries the rounding of partial sums of magnitude ~10, so its own magnitude says
    nothing about the attainable agreement (the CPU's sequential float sum is itself that far from
    the exact sum).  Bound: 1e-5 * max(|want|, 1, l1 mass) when the mass is given, else
    1e-5 * max(|want|, 1)."""
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    scale = np.maximum(np.abs(want), 1.0)
    if l1 is not None:
        scale = np.maximum(scale, np.asarray(l1, dtype=np.float64))
        if not score_tight(got, want, l1):     # the attainable bound, ten times tighter (see score_tight)
            return False
    return bool(np.all(np.abs(got - want) <= 1e-5 * scale))


def score_tight(got, want, l1):
    """What a re-ordered float32 sum actually achieves, asserted next to the stated tolerance so that the
    looser reading above cannot hide a real defect: sequential f32 summation of T terms carries an rms
    rounding error of about 0.5 * 2^-24 * sigma * T, i.e. ~3e-8 of the L1 mass whatever T is, so two
    orderings differ by ~4e-8 * mass rms and ~2.5e-7 * mass at the far tail of 10^7 rows.  Bound:
    1e-6 * max(l1 mass, 1) -- ten times tighter than 1e-5 * mass."""
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    scale = np.maximum(np.asarray(l1, dtype=np.float64), 1.0)
    return bool(np.all(np.abs(got - want) <= 1e-6 * scale))


def abs_leaf_forest(nodes):
    """The same trees with every LEAF value replaced by its magnitude: a row reaches the same leaves, so
    its score under this forest is the L1 mass of its summands (used to bound re-ordered sums on all rows,
    on the device, without 4 bytes x rows x trees of leaf ids).  Works for dense and sparse node arrays."""
    out = nodes.copy()
    leaf = out["bits"] < 0
    out["val"][leaf] = np.abs(out["val"][leaf])
    return out


def l1_mass_sparse(nodes, tree_offsets, leaf_ids):
    """sum_t |leaf value reached in tree t| per row from the oracle's leaf ids (sparse forests: node index in the tree)."""
    offs = np.asarray(tree_offsets, dtype=np.int64)[:-1][None, :]
    return np.abs(nodes["val"][offs + leaf_ids].astype(np.float64)).sum(axis=1)
