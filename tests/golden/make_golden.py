"""Generate tests/golden/*.npz by running the REFERENCE's own CPU code.

Runs only where /root/reference exists (the authoring container): it builds
oracle/_ref/libtahoe_ref.so (oracle/ref_shim.cu, which #includes the reference's
BaseTahoeTest.h in place) and records what the reference's loaders,
infer_one_tree loop and packer produce on seeded synthetic inputs.  The vectors
are committed; this script is the provenance.

    python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle import RefShim  # noqa: E402
from tahoe_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
MISSING = -999.0

# name: (trees, depth, cols, rows, leaf_prob, miss_frac, nan_frac, grid, store_inputs)
CASES = {
    "tiny":        (3, 2, 4, 16, 0.3, 0.2, 0.05, 0, True),
    "char_w":      (37, 5, 20, 400, 0.2, 0.02, 0.01, 0, True),     # max_fid < 32   -> 1-byte meta
    "short_w":     (64, 4, 300, 128, 0.2, 0.02, 0.01, 0, True),    # max_fid < 8192 -> 2-byte meta
    "int_w":       (16, 3, 9000, 32, 0.1, 0.01, 0.0, 0, True),     # wide           -> 4-byte meta
    "ties":        (40, 6, 12, 512, 0.1, 0.05, 0.02, 16, True),    # values/thresholds on a 1/16 grid
    "one_feature": (9, 4, 1, 64, 0.1, 0.1, 0.0, 0, True),          # max_fid == 0
    "depth0":      (5, 0, 3, 8, 0.0, 0.0, 0.0, 0, True),           # every tree is a single leaf
    "susy_like":   (500, 10, 18, 256, 0.0, 0.0, 0.0, 0, False),    # C2 forest shape, seed only
    "higgs_like":  (1000, 12, 28, 64, 0.0, 0.0, 0.0, 0, False),    # C3 forest shape, seed only
    "svhn_like":   (100, 8, 3072, 64, 0.0, 0.0, 0.0, 0, False),    # C1 forest shape, seed only
}
TRANSFORMS = [(0x0, 0.0, 0.0), (0x1, 0.0, 0.0), (0x11, 0.0, 0.25), (0x110, 0.5, -0.1),
              (0x111, 0.45, 0.0), (0x0, 0.0, 1.5)]


def main():
    ref = RefShim()
    for i, (name, (T, d, C, R, lp, mf, nf, grid, store)) in enumerate(CASES.items()):
        fseed, dseed = 7000 + i, 8000 + i
        nodes = synth.gen_dense_forest(fseed, T, d, C, leaf_prob=lp, grid=grid)
        data = synth.gen_data(dseed, R, C, miss_frac=mf, nan_frac=nf, missing=MISSING, grid=grid)
        out = dict(trees=T, depth=d, cols=C, rows=R, leaf_prob=lp, miss_frac=mf, nan_frac=nf,
                   grid=grid, forest_seed=fseed, data_seed=dseed, missing=np.float32(MISSING),
                   nodes_sha256=hashlib.sha256(nodes.tobytes()).hexdigest(),
                   data_sha256=hashlib.sha256(data.tobytes()).hexdigest())
        if store:
            out["nodes"] = nodes
            out["data"] = data
        for (flags, thr, bias) in TRANSFORMS:
            out[f"preds_{flags:03x}_{thr}_{bias}"] = ref.predict(
                nodes, T, d, data, MISSING, output=flags, threshold=thr, global_bias=bias)
        if store and T * synth.tree_num_nodes(d) <= 200_000 and d >= 1:
            bl, vals, bits = ref.pack(nodes, T, d, C)
            out["pack_bits_length"] = bl
            out["pack_values"] = vals
            out["pack_bits"] = bits
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
        print(name, "written")

    # text fixtures: files + what the reference's loaders parse from them
    nodes = synth.gen_dense_forest(7100, 3, 2, 4, leaf_prob=0.3)
    data = synth.gen_data(8100, 6, 4, miss_frac=0.2, missing=MISSING)
    mpath, dpath = os.path.join(HERE, "tiny_model.txt"), os.path.join(HERE, "tiny_data.txt")
    synth.write_model_text(mpath, nodes, 3, 2)
    synth.write_data_text(dpath, data, MISSING)
    # a hand-written file with the parser's corner cases: blank/garbage lines,
    # exponent notation, trailing text, CRLF
    with open(os.path.join(HERE, "quirks_model.txt"), "w") as f:
        f.write("2\n2 levels\n")
        rows = [("1", "0.5", "1", "1.0", "0"), ("0", "1e-1", "0", "0.25", "1"),
                ("3 trailing", "-2.5E0", "1", ".75", "1\r"),
                ("2", "abc", "0", "1", "0"), ("", "  7.25", "1", "0.5", "1"),
                ("1", "-0.0", "0", "0.5", "1")]
        for r in rows:
            f.write("\n".join(r) + "\n")
    with open(os.path.join(HERE, "quirks_data.txt"), "w") as f:
        f.write("3\n4\n-999.0\n")
        vals = ["0.5", "1e-1", "-999", "nan", "-999.0000001", "3.5x", "", "inf",
                "-inf", "  2", "0x10", "-0.0"]
        f.write("\n".join(vals) + "\n")
    rn, rt, rd = ref.load_model(mpath)
    rx, rm = ref.load_data(dpath)
    qn, qt, qd = ref.load_model(os.path.join(HERE, "quirks_model.txt"))
    qx, qm = ref.load_data(os.path.join(HERE, "quirks_data.txt"))
    np.savez_compressed(os.path.join(HERE, "text.npz"),
                        tiny_nodes=rn, tiny_trees=rt, tiny_depth=rd, tiny_data=rx, tiny_missing=rm,
                        tiny_preds=ref.predict(rn, rt, rd, rx, rm),
                        quirks_nodes=qn, quirks_trees=qt, quirks_depth=qd, quirks_data=qx,
                        quirks_missing=qm, quirks_preds=ref.predict(qn, qt, qd, qx, qm))
    print("text written")


if __name__ == "__main__":
    main()
