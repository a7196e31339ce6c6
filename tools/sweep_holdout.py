"""Out-of-sample check of the performance model (tahoe_select_strategy): forests the model's constants were
NOT fitted on, every applicable strategy timed at R = 1 ... 10^7, model pick vs measured best.

    python tools/sweep_holdout.py [max_rows] > profiles/r02_sweep_holdout.json

Held-out shapes (the constants in csrc/abi.cu `namespace model` came from configs 1-4): 200 x d6 x 64 cols,
50 x d14 x 28, 3000 x d8 x 18, two sparse shapes (300 trees <= 12 levels x 30 cols; 1000 trees <= 16 levels x
16 cols) and one importer-produced forest (a scikit-learn random forest: 100 trees, depth <= 12, 40 cols)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from sweep import time_strategy  # noqa: E402
from tahoe_b200 import api, synth  # noqa: E402


def forests():
    for name, T, d, C in (("dense_200x6x64", 200, 6, 64), ("dense_50x14x28", 50, 14, 28), ("dense_3000x8x18", 3000, 8, 18)):
        nodes = synth.gen_dense_forest(7000 + T, T, d, C)
        yield name, api.Forest.from_dense(nodes, T, d, C, missing=-999.0), C
    for name, T, D, C in (("sparse_300x12x30", 300, 12, 30), ("sparse_1000x16x16", 1000, 16, 16)):
        sp, offs, w, _ = synth.gen_sparse_forest(8000 + T, T, D, C)
        yield name, api.Forest.from_sparse(sp, offs, C, weights=w, missing=-999.0), C
    try:
        from sklearn.ensemble import RandomForestRegressor
        from tahoe_b200 import importers
        rng = np.random.default_rng(9)
        X = rng.uniform(-1, 1, size=(4000, 40)).astype(np.float32)
        y = np.sin(3 * X[:, 0]) + X[:, 1] * X[:, 2] + 0.1 * rng.normal(size=4000)
        m = RandomForestRegressor(n_estimators=100, max_depth=12, random_state=0, n_jobs=-1).fit(X, y)
        yield "sklearn_rf_100x12x40", importers.from_sklearn(m).forest(), 40
    except ImportError:
        pass


def main():
    max_rows = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
    sizes = [r for r in (1, 100, 10_000, 100_000, 1_000_000, 10_000_000) if r <= max_rows]
    out = {"forests": [], "sizes": sizes}
    hits = n = 0
    worst = 1.0
    for name, f, C in forests():
        rows_cap = min(max_rows, int(6e9 // (C * 4)))          # <= 6 GB of rows
        big = api.gen_data_device(4242, rows_cap, C)
        pbig = torch.empty(rows_cap, dtype=torch.float32, device="cuda")
        rec = {"name": name, "points": []}
        for R in sizes:
            if R > rows_cap:
                continue
            d, p = big[:R], pbig[:R]
            times = {api.STRATEGY_NAMES[s]: time_strategy(f, d, p, s) for s in api.ALL_STRATEGIES if f.applicable(s, R)}
            pick = api.STRATEGY_NAMES[f.select(R)[0]]
            best = min(times, key=times.get)
            slow = times[pick] / times[best]
            hits += pick == best; n += 1; worst = max(worst, slow)
            rec["points"].append({"rows": R, "ms": times, "best": best, "model_pick": pick, "pick_slowdown": slow})
            print(f"# {name:22s} R={R:>9} best={best:24s} pick={pick:24s} slowdown={slow:.2f} " +
                  " ".join(f"{k}={v:.4f}" for k, v in times.items()), file=sys.stderr, flush=True)
        rec["hit_rate"] = sum(q["model_pick"] == q["best"] for q in rec["points"]) / max(1, len(rec["points"]))
        rec["worst_pick_slowdown"] = max([q["pick_slowdown"] for q in rec["points"]], default=1.0)
        out["forests"].append(rec)
        del big, pbig
        f.close()
        torch.cuda.empty_cache()
    out["held_out_hit_rate"] = hits / max(1, n)
    out["held_out_worst_pick_slowdown"] = worst
    print(json.dumps(out))


if __name__ == "__main__":
    main()
