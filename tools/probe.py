"""Quick per-strategy timing probe (not the bench): python tools/probe.py [C2|C3|C1] [rows]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tahoe_b200 import api, synth

def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
    c = synth.CONFIGS[cfg]
    rows = int(sys.argv[2]) if len(sys.argv) > 2 else c["rows"]
    only = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else api.ALL_STRATEGIES
    t0 = time.time()
    if c["sparse"]:
        sp, offs, w, lv = synth.gen_sparse_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
        f = api.Forest.from_sparse(sp, offs, c["cols"], weights=w, missing=-999.0)
    else:
        nodes = synth.gen_dense_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
        f = api.Forest.from_dense(nodes, c["trees"], c["depth"], c["cols"], missing=-999.0)
    inf = f.info()
    print(f"{cfg}: pack {time.time()-t0:.1f}s nodes={inf.num_nodes} mean_depth={inf.mean_depth:.2f} chunks={inf.num_chunks} tpc={inf.trees_per_chunk} rank_bits={inf.rank_bits} quant={inf.quantized} col_shift={inf.col_shift}", flush=True)
    d = api.gen_data_device(synth.DATA_SEED[cfg], rows, c["cols"])
    p = torch.empty(rows, dtype=torch.float32, device="cuda")
    ref = None
    for s in only:
        if not f.applicable(s, rows):
            print(api.STRATEGY_NAMES[s], "not applicable"); continue
        n_w, n_t = (1, 2) if s in (api.DIRECT, api.SHARED_DATA_PER_THREAD, api.SHARED_DATA) else (3, 5)
        for _ in range(n_w): f.predict(d, p, strategy=s)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_t): f.predict(d, p, strategy=s)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n_t
        out = p.cpu().numpy()
        if ref is None: ref = out.copy()
        print(f"{api.STRATEGY_NAMES[s]:24s} {ms:9.3f} ms  {rows/ms/1e3:9.2f} Msamples/s  maxdiff_vs_first={np.abs(out-ref).max():.3g}", flush=True)
    print("model pick:", f.select(rows))

if __name__ == "__main__":
    main()
