"""Batch-size sweep (BASELINE.json config 5): time every applicable strategy at every batch size and
compare the performance model's pick with the measured best (exhaustive enumeration).

    python tools/sweep.py C2 [max_rows] > profiles/rNN_sweep_C2.json
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from tahoe_b200 import api, synth  # noqa: E402


def time_strategy(f, d, p, s, budget_s=0.25):
    """median-free: warm up once, then as many predicts as fit the budget (3..50), events on the stream"""
    f.predict(d, p, strategy=s)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); f.predict(d, p, strategy=s); e1.record(); torch.cuda.synchronize()
    one = max(e0.elapsed_time(e1) * 1e-3, 1e-6)
    n = int(min(50, max(3, budget_s / one)))
    best = float("inf")
    for _ in range(3 if one < 0.05 else 1):
        e0.record()
        for _ in range(n):
            f.predict(d, p, strategy=s)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / n)
    return best


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
    c = synth.CONFIGS[cfg]
    max_rows = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000_000
    if c["sparse"]:
        sp, offs, w, lv = synth.gen_sparse_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
        f = api.Forest.from_sparse(sp, offs, c["cols"], weights=w, missing=-999.0)
    else:
        nodes = synth.gen_dense_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
        f = api.Forest.from_dense(nodes, c["trees"], c["depth"], c["cols"], missing=-999.0)
    sizes = []
    k = 1
    while k <= max_rows:
        sizes.append(k)
        if 3 * k <= max_rows:
            sizes.append(3 * k)
        k *= 10
    big = api.gen_data_device(synth.DATA_SEED[cfg], max(sizes), c["cols"])
    pbig = torch.empty(max(sizes), dtype=torch.float32, device="cuda")
    out = {"config": cfg, "shape": c, "points": []}
    hits, worst = 0, 1.0
    for R in sizes:
        d, p = big[:R], pbig[:R]
        times = {}
        for s in api.ALL_STRATEGIES:
            if f.applicable(s, R):
                times[api.STRATEGY_NAMES[s]] = time_strategy(f, d, p, s)
        pick, pred = f.select(R)
        best = min(times, key=times.get)
        pick_name = api.STRATEGY_NAMES[pick]
        slowdown = times[pick_name] / times[best]
        hits += pick_name == best
        worst = max(worst, slowdown)
        out["points"].append({"rows": R, "ms": times, "best": best, "model_pick": pick_name,
                              "model_ms": {api.STRATEGY_NAMES[i]: (t * 1e3 if t < 1e30 else None) for i, t in enumerate(pred)},
                              "pick_slowdown": slowdown})
        print(f"# R={R:>10} best={best:24s} pick={pick_name:24s} slowdown={slowdown:.2f} " +
              " ".join(f"{k}={v:.4f}" for k, v in times.items()), file=sys.stderr, flush=True)
    out["hit_rate"] = hits / len(sizes)
    out["worst_pick_slowdown"] = worst
    print(json.dumps(out))


if __name__ == "__main__":
    main()
