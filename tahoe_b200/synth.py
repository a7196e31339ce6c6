"""Seeded synthetic forests and sample matrices.

The reference's models and datasets are Google-Drive downloads (reference
README.md:25-30) and its own generators are unusable (generate_forest passes its
arguments in the wrong order, BaseTahoeTest.h:228-229 vs :132-135) and never
called (:81, :84).  This module is the replacement: a counter-based RNG
(splitmix64 keyed by (seed, stream, index)) so that numpy on the host and the
CUDA generator in csrc/gen.cu produce bit-identical values without copying.

Value law follows the reference's U(-1,1) (BaseTahoeTest.h:123-130).

Node encodings are the reference's (Struct.h:44-59, 103-108):
    dense_node_t  {float weight; float val; int bits}   bits = fid | def_left<<30 | is_leaf<<31
    sparse_node_t {float val; int bits; int left_idx}
"""
from __future__ import annotations

import numpy as np

DENSE_NODE = np.dtype([("weight", "<f4"), ("val", "<f4"), ("bits", "<i4")])
SPARSE_NODE = np.dtype([("val", "<f4"), ("bits", "<i4"), ("left_idx", "<i4")])

FID_MASK = (1 << 30) - 1
DEF_LEFT_MASK = 1 << 30
IS_LEAF_MASK = -(1 << 31)  # int32 sign bit

OUT_RAW, OUT_AVG, OUT_SIGMOID, OUT_THRESHOLD = 0x0, 0x1, 0x10, 0x100

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
_GOLD = np.uint64(0x9E3779B97F4A7C15)
_MIX1 = np.uint64(0xBF58476D1CE4E5B9)
_MIX2 = np.uint64(0x94D049BB133111EB)
_STREAM = np.uint64(0x632BE59BD9B4E019)

# RNG streams (shared with csrc/gen.cu)
S_DATA = 1
S_FID, S_VAL, S_DEF, S_SPLIT, S_LEAF = 11, 12, 13, 14, 15


def splitmix64(x: np.ndarray) -> np.ndarray:
    """One splitmix64 output step on uint64 arrays (wrap-around arithmetic)."""
    with np.errstate(over="ignore"):
        z = (x + _GOLD).astype(np.uint64)
        z = ((z ^ (z >> np.uint64(30))) * _MIX1).astype(np.uint64)
        z = ((z ^ (z >> np.uint64(27))) * _MIX2).astype(np.uint64)
        return z ^ (z >> np.uint64(31))


def rnd64(seed: int, stream: int, idx: np.ndarray) -> np.ndarray:
    """r = splitmix64(splitmix64(seed + STREAM*(stream+1)) ^ idx)."""
    with np.errstate(over="ignore"):
        key = splitmix64(np.array([(seed + int(_STREAM) * (stream + 1)) & 0xFFFFFFFFFFFFFFFF],
                                  dtype=np.uint64))[0]
    return splitmix64(np.asarray(idx, dtype=np.uint64) ^ key)


def u24_to_unit(u24: np.ndarray) -> np.ndarray:
    """24-bit integer -> float32 in [-1, 1), exact (same expression as gen.cu)."""
    return u24.astype(np.float32) * np.float32(1.0 / 8388608.0) - np.float32(1.0)


def gen_data(seed: int, rows: int, cols: int, *, miss_frac: float = 0.0,
             nan_frac: float = 0.0, missing: float = -999.0, grid: int = 0,
             row0: int = 0) -> np.ndarray:
    """Row-major float32 [rows, cols]; element index = (row0+row)*cols + col.

    top 24 bits of r -> value; low 24 bits -> missing coin; bits 24..39 -> NaN coin.
    grid > 0 snaps values to multiples of 1/grid (forces threshold ties in tests).
    """
    idx = (np.arange(rows * cols, dtype=np.uint64) + np.uint64(row0 * cols))
    r = rnd64(seed, S_DATA, idx)
    v = u24_to_unit(r >> np.uint64(40))
    if grid:
        v = (np.round(v * np.float32(grid)) / np.float32(grid)).astype(np.float32)
    if nan_frac > 0.0:
        coin = ((r >> np.uint64(24)) & np.uint64(0xFFFF)).astype(np.int64)
        v[coin < int(nan_frac * 65536.0)] = np.float32(np.nan)
    if miss_frac > 0.0:
        coin = (r & np.uint64(0xFFFFFF)).astype(np.int64)
        v[coin < int(miss_frac * 16777216.0)] = np.float32(missing)
    return v.reshape(rows, cols)


def tree_num_nodes(depth: int) -> int:
    return (1 << (depth + 1)) - 1


def gen_dense_forest(seed: int, num_trees: int, depth: int, num_cols: int, *,
                     leaf_prob: float = 0.0, grid: int = 0,
                     def_left_mode: str = "random") -> np.ndarray:
    """Dense forest in the reference's AoS layout: [num_trees * (2^(depth+1)-1)] DENSE_NODE.

    Heap order per tree (root 0, children 2k+1 / 2k+2).  Bottom level is always
    leaf (the reference never bounds `curr`, BaseTahoeTest.h:446-454).  Above it a
    node whose parent is internal becomes a leaf with probability leaf_prob; slots
    below an early leaf are unreachable and are filled with leaf-flagged noise.
    weight = cover fraction: root 1, left child w*parent, right (1-w)*parent,
    w ~ U(0.05, 0.95).
    """
    npt = tree_num_nodes(depth)
    n = num_trees * npt
    idx = np.arange(n, dtype=np.uint64)
    fid = (rnd64(seed, S_FID, idx) >> np.uint64(33)).astype(np.int64) % max(num_cols, 1)
    val = u24_to_unit(rnd64(seed, S_VAL, idx) >> np.uint64(40))
    if grid:
        val = (np.round(val * np.float32(grid)) / np.float32(grid)).astype(np.float32)
    if def_left_mode == "random":
        dl = (rnd64(seed, S_DEF, idx) >> np.uint64(63)).astype(bool)
    elif def_left_mode == "all":
        dl = np.ones(n, dtype=bool)
    else:
        dl = np.zeros(n, dtype=bool)
    split = (0.05 + 0.9 * ((rnd64(seed, S_SPLIT, idx) >> np.uint64(40)).astype(np.float64)
                           / 16777216.0)).astype(np.float32)
    coin = (rnd64(seed, S_LEAF, idx) >> np.uint64(40)).astype(np.float64) / 16777216.0

    node = np.arange(npt)
    level = np.floor(np.log2(node + 1)).astype(np.int64)
    is_leaf = np.zeros((num_trees, npt), dtype=bool)
    weight = np.ones((num_trees, npt), dtype=np.float32)
    coin2 = coin.reshape(num_trees, npt)
    split2 = split.reshape(num_trees, npt)
    for lv in range(depth + 1):
        cur = node[level == lv]
        if lv == depth:
            is_leaf[:, cur] = True
        else:
            if lv > 0:
                is_leaf[:, cur] |= coin2[:, cur] < leaf_prob
            else:
                is_leaf[:, cur] = False if depth > 0 else True
        if lv > 0:
            par = (cur - 1) // 2
            is_leaf[:, cur] |= is_leaf[:, par]
            left = (cur % 2) == 1
            w = np.where(left[None, :], split2[:, par], np.float32(1.0) - split2[:, par])
            weight[:, cur] = (weight[:, par] * w).astype(np.float32)
    out = np.zeros(n, dtype=DENSE_NODE)
    out["weight"] = weight.reshape(-1)
    out["val"] = val
    bits = fid.astype(np.int64) & FID_MASK
    bits |= np.where(dl, DEF_LEFT_MASK, 0)
    bits = bits.astype(np.int64)
    leaf_flat = is_leaf.reshape(-1)
    bits = np.where(leaf_flat, bits | (1 << 31), bits)
    out["bits"] = (bits & 0xFFFFFFFF).astype(np.uint32).view(np.int32)
    return out


def dense_to_sparse(nodes: np.ndarray, num_trees: int, depth: int):
    """Compact the reachable nodes of a dense forest, breadth-first per tree,
    children stored as an adjacent pair (left_idx relative to the tree's first
    node; Struct.h:2244 addressing).  Returns (sparse_nodes, tree_offsets[T+1],
    weights, dense_ids) where dense_ids maps every compact node to its dense
    heap index (used to compare leaf ids between the two layouts)."""
    npt = tree_num_nodes(depth)
    bits = nodes["bits"].reshape(num_trees, npt)
    leaf = bits < 0
    sp_chunks, w_chunks, id_chunks, offsets = [], [], [], [0]
    for t in range(num_trees):
        order = [0]
        left_idx = []
        head = 0
        while head < len(order):
            k = order[head]
            if leaf[t, k]:
                left_idx.append(0)
            else:
                left_idx.append(len(order))
                order.append(2 * k + 1)
                order.append(2 * k + 2)
            head += 1
        order = np.asarray(order, dtype=np.int64)
        sp = np.zeros(len(order), dtype=SPARSE_NODE)
        src = nodes[t * npt + order]
        sp["val"] = src["val"]
        sp["bits"] = src["bits"]
        sp["left_idx"] = np.asarray(left_idx, dtype=np.int32)
        sp_chunks.append(sp)
        w_chunks.append(src["weight"].copy())
        id_chunks.append(order.astype(np.int32))
        offsets.append(offsets[-1] + len(order))
    return (np.concatenate(sp_chunks), np.asarray(offsets, dtype=np.int64),
            np.concatenate(w_chunks), np.concatenate(id_chunks))


def gen_sparse_forest(seed: int, num_trees: int, max_depth: int, num_cols: int, *,
                      leaf_prob: float = 0.30, spine: bool = True):
    """Deep unbalanced forest generated directly in compact form (config C4,
    SURVEY.md 8d): every tree has one forced root-to-max_depth spine; every other
    node at levels 1..max_depth-1 is a leaf with probability leaf_prob; level
    max_depth nodes are always leaves.  Breadth-first storage per tree, children
    adjacent.  Vectorised level by level across all trees.

    Returns (sparse_nodes, tree_offsets[T+1], weights, levels)."""
    # frontier state per live node: tree id, path code (for rng index), on_spine, weight
    tree = np.arange(num_trees, dtype=np.int64)
    code = np.ones(num_trees, dtype=np.uint64)            # heap path code 1 = root (1-based)
    on_spine = np.ones(num_trees, dtype=bool) if spine else np.zeros(num_trees, dtype=bool)
    weight = np.ones(num_trees, dtype=np.float32)
    per_level = []
    for lv in range(max_depth + 1):
        m = tree.shape[0]
        ridx = (tree.astype(np.uint64) << np.uint64(40)) ^ (code * np.uint64(0x2545F4914F6CDD1D))
        fid = (rnd64(seed, S_FID, ridx) >> np.uint64(33)).astype(np.int64) % max(num_cols, 1)
        val = u24_to_unit(rnd64(seed, S_VAL, ridx) >> np.uint64(40))
        dl = (rnd64(seed, S_DEF, ridx) >> np.uint64(63)).astype(bool)
        split = (0.05 + 0.9 * ((rnd64(seed, S_SPLIT, ridx) >> np.uint64(40)).astype(np.float64)
                               / 16777216.0)).astype(np.float32)
        coin = (rnd64(seed, S_LEAF, ridx) >> np.uint64(40)).astype(np.float64) / 16777216.0
        if lv == max_depth:
            is_leaf = np.ones(m, dtype=bool)
        elif lv == 0:
            is_leaf = np.zeros(m, dtype=bool)
        else:
            is_leaf = (coin < leaf_prob) & ~on_spine
        per_level.append(dict(tree=tree, fid=fid, val=val, dl=dl, is_leaf=is_leaf,
                              weight=weight, level=np.full(m, lv, dtype=np.int32)))
        if lv == max_depth:
            break
        keep = ~is_leaf
        ptree, pcode, pspine, pw, psplit = tree[keep], code[keep], on_spine[keep], weight[keep], split[keep]
        # spine continues through the child chosen by one more coin
        go_right = (rnd64(seed, S_LEAF + 1, (ptree.astype(np.uint64) << np.uint64(40)) ^ pcode)
                    >> np.uint64(63)).astype(bool)
        k = ptree.shape[0]
        tree = np.repeat(ptree, 2)
        code = np.empty(2 * k, dtype=np.uint64)
        code[0::2] = pcode * np.uint64(2)
        code[1::2] = pcode * np.uint64(2) + np.uint64(1)
        on_spine = np.zeros(2 * k, dtype=bool)
        on_spine[0::2] = pspine & ~go_right
        on_spine[1::2] = pspine & go_right
        weight = np.empty(2 * k, dtype=np.float32)
        weight[0::2] = pw * psplit
        weight[1::2] = pw * (np.float32(1.0) - psplit)
    # assemble breadth-first per tree: sort by (tree, level) keeping in-level order
    allt = np.concatenate([p["tree"] for p in per_level])
    alll = np.concatenate([p["level"] for p in per_level])
    pos = np.concatenate([np.arange(p["tree"].shape[0], dtype=np.int64) for p in per_level])
    order = np.lexsort((pos, alll, allt))
    cat = {k: np.concatenate([p[k] for p in per_level])[order]
           for k in ("tree", "fid", "val", "dl", "is_leaf", "weight", "level")}
    n = cat["tree"].shape[0]
    counts = np.bincount(cat["tree"], minlength=num_trees)
    offsets = np.zeros(num_trees + 1, dtype=np.int64)
    np.cumsum(counts, out=offsets[1:])
    # left_idx: within a tree in BFS order, the j-th internal node's children sit at 1+2j, 2+2j
    internal = ~cat["is_leaf"]
    rank_in_tree = np.cumsum(internal) - 1
    first_rank = np.zeros(num_trees, dtype=np.int64)
    tree_first_internal_rank = np.concatenate([[0], np.cumsum(np.bincount(cat["tree"][internal],
                                                                        minlength=num_trees))[:-1]])
    first_rank[:] = tree_first_internal_rank
    j = rank_in_tree - first_rank[cat["tree"]]
    left_idx = np.where(internal, 1 + 2 * j, 0).astype(np.int32)
    sp = np.zeros(n, dtype=SPARSE_NODE)
    sp["val"] = cat["val"]
    bits = cat["fid"].astype(np.int64) & FID_MASK
    bits |= np.where(cat["dl"], DEF_LEFT_MASK, 0)
    bits = np.where(cat["is_leaf"], bits | (1 << 31), bits)
    sp["bits"] = (bits & 0xFFFFFFFF).astype(np.uint32).view(np.int32)
    sp["left_idx"] = left_idx
    return sp, offsets, cat["weight"].astype(np.float32), cat["level"]


def write_model_text(path: str, nodes: np.ndarray, num_trees: int, depth: int) -> None:
    """Reference model text grammar (BaseTahoeTest.h:279-315)."""
    bits = nodes["bits"].astype(np.int64)
    fid = bits & FID_MASK
    dl = (bits & DEF_LEFT_MASK) != 0
    leaf = nodes["bits"] < 0
    with open(path, "w") as f:
        f.write(f"{num_trees}\n{depth + 1}\n")
        lines = []
        for i in range(nodes.shape[0]):
            lines.append(f"{int(fid[i])}\n{float(nodes['val'][i])!r}\n{int(dl[i])}\n"
                         f"{float(nodes['weight'][i])!r}\n{int(leaf[i])}\n")
        f.write("".join(lines))


def write_data_text(path: str, data: np.ndarray, missing: float) -> None:
    """Reference data text grammar (BaseTahoeTest.h:366-388)."""
    rows, cols = data.shape
    with open(path, "w") as f:
        f.write(f"{rows}\n{cols}\n{float(missing)!r}\n")
        f.write("".join(f"{float(v)!r}\n" for v in data.reshape(-1)))


# The BASELINE.json configurations.
CONFIGS = {
    "C1": dict(name="SVHN-shaped", rows=50_000, cols=3072, trees=100, depth=8, sparse=False),
    "C2": dict(name="SUSY-shaped", rows=5_000_000, cols=18, trees=500, depth=10, sparse=False),
    "C3": dict(name="Higgs-shaped", rows=10_000_000, cols=28, trees=1000, depth=12, sparse=False),
    "C4": dict(name="deep-unbalanced", rows=1_000_000, cols=50, trees=2000, depth=20, sparse=True),
}
FOREST_SEED = {"C1": 20240002, "C2": 20240003, "C3": 20240004, "C4": 20240005}
DATA_SEED = {"C1": 20240102, "C2": 20240103, "C3": 20240104, "C4": 20240105}
