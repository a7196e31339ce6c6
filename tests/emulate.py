"""Test-only numpy emulators of what each kernel does with each PACKED layout.
They let the CPU suite localise packer bugs without a GPU.  Not a product path."""
import numpy as np

LEAF, DEFLEFT, EXCH, FIDMASK = 1 << 31, 1 << 30, 1 << 29, (1 << 29) - 1
SD_MISSRIGHT, SD_FLIP, SD_LEAF, SD_OFFMASK = 1 << 31, 1 << 30, 1 << 29, (1 << 29) - 1     # B32 blocks (SHARED_DATA)


def _right(v, thr, def_left, missing):
    with np.errstate(invalid="ignore"):
        miss = np.abs((v - np.float32(missing)).astype(np.float32)) <= np.float32(1e-6)
        return np.where(miss, ~def_left, v >= thr)


def heap_traverse(heap, T, depth, data, missing):
    """H8: uint32 pairs [T][2^(d+1)][2], 1-based heap.  -> (leaf values [R,T], leaf ids [R,T])"""
    W = 1 << (depth + 1)
    h = heap.reshape(T, W, 2)
    R = data.shape[0]
    vals = np.zeros((R, T), np.float32)
    ids = np.zeros((R, T), np.int32)
    for t in range(T):
        n = np.ones(R, np.int64)
        for _ in range(depth + 1):
            e = h[t, n]
            leaf = (e[:, 1] & LEAF) != 0
            fid = (e[:, 1] & 0x3FFFFFFF).astype(np.int64)
            fid = np.where(leaf, 0, fid)
            v = data[np.arange(R), fid]
            c = _right(v, e[:, 0].view(np.float32), (e[:, 1] & DEFLEFT) != 0, missing)
            n = np.where(leaf, n, 2 * n + c)
        e = h[t, n]
        assert np.all(e[:, 1] & LEAF)
        vals[:, t] = e[:, 0].view(np.float32)
        ids[:, t] = n - 1
    return vals, ids


def interleaved_traverse(inter, orig_ids, tree_order, tpad, depth, T, data, missing):
    """I8: [N][tpad][2] with exchange bits.  -> (values [R,T], ids [R,T]) in ORIGINAL tree order"""
    N = (1 << (depth + 1)) - 1
    h = inter.reshape(N, tpad, 2)
    oid = orig_ids.reshape(N, tpad)
    R = data.shape[0]
    vals = np.zeros((R, T), np.float32)
    ids = np.zeros((R, T), np.int32)
    for slot in range(tpad):
        t = tree_order[slot]
        n = np.zeros(R, np.int64)
        for _ in range(depth + 1):
            e = h[n, slot]
            leaf = (e[:, 1] & LEAF) != 0
            fid = np.where(leaf, 0, (e[:, 1] & FIDMASK).astype(np.int64))
            v = data[np.arange(R), fid]
            c = _right(v, e[:, 0].view(np.float32), (e[:, 1] & DEFLEFT) != 0, missing)
            c = np.where((e[:, 1] & EXCH) != 0, ~c, c)
            n = np.where(leaf, n, 2 * n + 1 + c)
        e = h[n, slot]
        assert np.all(e[:, 1] & LEAF)
        if t < 0:
            assert np.all(e[:, 0] == 0)
            continue
        vals[:, t] = e[:, 0].view(np.float32)
        ids[:, t] = oid[n, slot]
    return vals, ids


def blocked_traverse(top_nodes, blocks, orig_bottom, tree_order, tpad, depth, K, T, data, missing):
    """B32 (SHARED_DATA): K top levels [2^K-1][tpad][2] with leaves pushed down, then two-level blocks
    [block][tpad][8] = {root, left, right, -}.  Exactly `depth` decisions, the kernel's index algebra.
    -> (values [R,T], ids [R,T]) in ORIGINAL tree order"""
    top = top_nodes.reshape((1 << K) - 1, tpad, 2)
    blk = blocks.reshape(-1, tpad, 8)
    ob = orig_bottom.reshape(1 << depth, tpad)
    J = (depth + 1 - K) // 2
    assert K >= 1 and J >= 1 and K + 2 * J == depth + 1
    R = data.shape[0]
    rows = np.arange(R)
    vals = np.zeros((R, T), np.float32)
    ids = np.zeros((R, T), np.int32)

    def decide(val, meta):
        # the walk's own node word (kSd* in forest.h): byte offset of the feature; bit 31 = a missing value takes
        # child slot 1; bit 30 = the children are exchanged (v >= thr takes slot 0); bit 29 = pushed-down leaf
        assert np.all((meta & SD_OFFMASK) % 4 == 0) and np.all(((meta & SD_LEAF) == 0) | ((meta & SD_OFFMASK) == 0))
        v = data[rows, (meta & SD_OFFMASK).astype(np.int64) // 4]
        with np.errstate(invalid="ignore"):
            miss = np.isnan(v) if np.isnan(missing) else np.abs((v - np.float32(missing)).astype(np.float32)) <= np.float32(1e-6)
            ge = v >= val.view(np.float32)
        return np.where(miss, (meta & SD_MISSRIGHT) != 0, ge != ((meta & SD_FLIP) != 0)).astype(np.int64)

    for slot in range(tpad):
        t = tree_order[slot]
        q = np.zeros(R, np.int64)
        for l in range(K):
            e = top[(1 << l) - 1 + q, slot]
            q = 2 * q + decide(e[:, 0], e[:, 1])
        base = 0
        for j in range(J):
            w = blk[base + q, slot]
            c0 = decide(w[:, 0], w[:, 1])
            cv = np.where(c0 == 1, w[:, 4], w[:, 2])
            cm = np.where(c0 == 1, w[:, 5], w[:, 3])
            if j == J - 1:
                q = 2 * q + c0
                assert np.all(cm & SD_LEAF)
                leaf_val = cv
            else:
                q = 4 * q + 2 * c0 + decide(cv, cm)
            base += 1 << (K + 2 * j)
        assert base == blk.shape[0]
        if t < 0:
            assert np.all(leaf_val == 0)
            continue
        vals[:, t] = leaf_val.view(np.float32)
        ids[:, t] = ob[q, slot]
    return vals, ids


def quantize_rows(data, table, offsets, rank_shift, miss_code, missing):
    """k_quantize: float rows -> rank words [R,C] (uint32)."""
    R, Ccols = data.shape
    out = np.zeros((R, Ccols), np.uint32)
    with np.errstate(invalid="ignore"):
        miss = np.abs((data - np.float32(missing)).astype(np.float32)) <= np.float32(1e-6)
    for c in range(Ccols):
        tab = table[offsets[c]:offsets[c + 1]]
        v = data[:, c]
        rank = np.searchsorted(tab, v, side="right")      # #{thr <= v}
        rank = np.where(np.isnan(v), 0, rank)
        out[:, c] = (rank.astype(np.uint64) << np.uint64(rank_shift)).astype(np.uint32)
    out[miss] = np.uint32(miss_code)
    return out


def quantize_rows_lut(data, table, offsets, lut, lut_min, lut_scale, rank_shift, miss_code, missing):
    """k_quantize exactly as the device does it: bucket LUT + search inside the bucket."""
    R, Ccols = data.shape
    B = lut.shape[0] // Ccols - 1
    lut = lut.reshape(Ccols, B + 1)
    out = np.zeros((R, Ccols), np.uint32)
    with np.errstate(invalid="ignore", over="ignore"):
        miss = np.abs((data - np.float32(missing)).astype(np.float32)) <= np.float32(1e-6)
        for c in range(Ccols):
            tab = table[offsets[c]:offsets[c + 1]]
            v = data[:, c]
            x = ((v - lut_min[c]).astype(np.float32) * lut_scale[c]).astype(np.float32)
            b = np.where(x >= np.float32(B), B - 1, np.trunc(np.nan_to_num(x, nan=0.0, posinf=B, neginf=-1)).astype(np.int64))
            b = np.clip(b, 0, B - 1)
            lo, hi = lut[c, b], lut[c, b + 1]
            rank = np.zeros(R, np.int64)
            for i in range(R):
                if v[i] == v[i]:
                    rank[i] = lo[i] + np.searchsorted(tab[lo[i]:hi[i]], v[i], side="right")
            out[:, c] = (rank.astype(np.uint64) << np.uint64(rank_shift)).astype(np.uint32)
    out[miss] = np.uint32(miss_code)
    return out


def quant_traverse(qwords, leaf_origin, T, depth, F, col_mask, miss_code, col_shift=7):
    """k_smem_forest: rank words + Q4 trees.  -> (values [R,T], ids [R,T])"""
    W, L = 1 << (depth + 1), 1 << depth
    q = qwords.reshape(T, W)
    lo = leaf_origin.reshape(T, L)
    R = F.shape[0]
    vals = np.zeros((R, T), np.float32)
    ids = np.zeros((R, T), np.int32)
    for t in range(T):
        n = np.ones(R, np.int64)
        for _ in range(depth):
            N = q[t, n]
            col = ((N & np.uint32(col_mask)) >> np.uint32(col_shift)).astype(np.int64)
            Fv = F[np.arange(R), col]
            c = Fv > N
            c = np.where(Fv == np.uint32(miss_code), (N & np.uint32(1)) == 0, c)
            n = 2 * n + c
        vals[:, t] = q[t, n].view(np.float32)
        ids[:, t] = lo[t, n - L]
    return vals, ids


def sparse_bfs_traverse(nodes8, orig, offsets, fid_bits, cut, data, missing):
    """T8: breadth-first 8-byte sparse nodes {val, leaf<<31 | def_left<<30 | delta<<fid_bits | fid}.
    -> (values [R,T], caller's leaf node index [R,T], walks that stepped past `cut` [R,T] bool)"""
    nd = nodes8.reshape(-1, 2)
    T = len(offsets) - 1
    R = data.shape[0]
    rows = np.arange(R)
    vals = np.zeros((R, T), np.float32)
    ids = np.zeros((R, T), np.int32)
    deep = np.zeros((R, T), bool)
    fmask = (1 << fid_bits) - 1
    for t in range(T):
        base = int(offsets[t])
        n = np.zeros(R, np.int64)
        while True:
            e = nd[base + n]
            leaf = (e[:, 1] & LEAF) != 0
            if leaf.all():
                break
            fid = np.where(leaf, 0, (e[:, 1] & fmask).astype(np.int64))
            c = _right(data[rows, fid], e[:, 0].view(np.float32), (e[:, 1] & DEFLEFT) != 0, missing)
            delta = ((e[:, 1] & 0x3FFFFFFF) >> fid_bits).astype(np.int64)
            assert np.all(delta[~leaf] >= 1)
            n = np.where(leaf, n, n + delta + c)
            deep[:, t] |= n >= cut
        vals[:, t] = nd[base + n, 0].view(np.float32)
        ids[:, t] = orig[base + n]
    return vals, ids, deep


CONT = 0xFFC00000


def qtop_traverse(qwords, leaf_origin, blocks, block_orig, block_offsets, T, depth, F, col_mask, miss_code, col_shift, tpc):
    """QT (k_smem_forest_deep): dense rank-quantised tops [T'][2^(depth+1)] + two-level blocks
    {root, left, right, -} x {rank word | leaf value, leaf<<31 | block delta} for the walks that leave
    them.  F = rank words [R,C].
    -> (values [R,T], caller's leaf node index [R,T], walks parked [R,T] bool, block loads [R,T])"""
    W, L = 1 << (depth + 1), 1 << depth
    q = qwords.reshape(-1, W)
    lo = leaf_origin.reshape(T, L)
    bl = blocks.reshape(-1, 8)
    bo = block_orig.reshape(-1, 3)
    R = F.shape[0]
    rows = np.arange(R)
    vals = np.zeros((R, T), np.float32)
    ids = np.zeros((R, T), np.int32)
    parked = np.zeros((R, T), bool)
    loads = np.zeros((R, T), np.int32)
    # the dummy block idle slots sit on: a leaf root worth +0.0f
    assert bl.shape[0] == block_offsets[T] + 1 and bl[-1, 0] == 0 and (bl[-1, 1] >> 31) == 1

    def right(Fv, N):
        c = Fv > N
        return np.where(Fv == np.uint32(miss_code), (N & np.uint32(1)) == 0, c)

    def feat(r, N):
        return F[r, ((N & np.uint32(col_mask)) >> np.uint32(col_shift)).astype(np.int64)]

    for t in range(T):
        n = np.ones(R, np.int64)
        for _ in range(depth):
            N = q[t, n]
            n = 2 * n + right(feat(rows, N), N)
        v = q[t, n]
        cont = (v & np.uint32(CONT)) == np.uint32(CONT)
        parked[:, t] = cont
        vals[~cont, t] = v[~cont].view(np.float32)
        ids[~cont, t] = lo[t, n[~cont] - L]
        assert np.all(lo[t, n[~cont] - L] >= 0)
        # the parked walks go on in the blocks; the marker's index is relative to the first block of
        # the chunk's first tree
        r = np.nonzero(cont)[0]
        base = int(block_offsets[t // tpc * tpc])
        b = base + (v[r] & np.uint32(~CONT & 0xFFFFFFFF)).astype(np.int64)
        first = True
        while r.size:
            assert np.all(b >= block_offsets[t]) and np.all(b < block_offsets[t + 1])
            loads[r, t] += 1
            e = bl[b]
            root_leaf = (e[:, 1] >> 31) != 0
            assert not (first and root_leaf.any())        # a marker only stands for an internal node
            first = False
            rx = np.where(root_leaf, np.uint32(0), e[:, 0])   # (a leaf's word is its float value, not a node)
            c0 = right(feat(r, rx), rx)
            cx = np.where(c0, e[:, 4], e[:, 2])
            cy = np.where(c0, e[:, 5], e[:, 3])
            child_leaf = (cy >> 31) != 0
            done = root_leaf | child_leaf
            vals[r[done], t] = np.where(root_leaf, e[:, 0], cx)[done].view(np.float32)
            pos = np.where(root_leaf, 0, np.where(c0, 2, 1))
            ids[r[done], t] = bo[b[done], pos[done]]
            assert np.all(bo[b[done], pos[done]] >= 0)
            cxm = np.where(done, np.uint32(0), cx)
            c1 = right(feat(r, cxm), cxm)
            b = b + cy.astype(np.int64) + c1
            r, b = r[~done], b[~done]
    return vals, ids, parked, loads
