"""Small batches (where SHARED_DATA is the pick on every config): per-call time of back-to-back predicts, wall clock
(host launch path + kernel, whichever is longer) -- python tools/probe_small_batch.py [CFG ...]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tahoe_b200 import api, synth
for cfg in (sys.argv[1:] or ["C1", "C2", "C3"]):
    c = synth.CONFIGS[cfg]
    nodes = synth.gen_dense_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
    f = api.Forest.from_dense(nodes, c["trees"], c["depth"], c["cols"], missing=-999.0)
    big = api.gen_data_device(synth.DATA_SEED[cfg], 1024, c["cols"])
    out = torch.empty(1024, dtype=torch.float32, device="cuda")
    for R in (1, 32, 1024):
        d, p = big[:R], out[:R]
        s = f.select(R)[0]
        for _ in range(20): f.predict(d, p, strategy=s)
        torch.cuda.synchronize()
        n = 2000
        t0 = time.perf_counter()
        for _ in range(n): f.predict(d, p, strategy=s)
        t_issue = time.perf_counter() - t0
        torch.cuda.synchronize()
        t_all = time.perf_counter() - t0
        # the C ABI called with ready-made pointers (what a C++ caller pays): no tensor checks, no stream lookup
        L, h = api.lib(), f._h
        dp, pp, st = d.data_ptr(), p.data_ptr(), torch.cuda.current_stream().cuda_stream
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n): L.tahoe_predict(h, s, dp, R, pp, None, st)
        t_raw_issue = time.perf_counter() - t0
        torch.cuda.synchronize()
        t_raw = time.perf_counter() - t0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f.predict(d, p, strategy=s); e1.record(); torch.cuda.synchronize()
        print(f"{cfg} R={R:5d} {api.STRATEGY_NAMES[s]:12s} host issue {t_issue/n*1e6:6.2f} us/call, back-to-back {t_all/n*1e6:6.2f} us/call | raw C ABI: issue {t_raw_issue/n*1e6:6.2f}, back-to-back {t_raw/n*1e6:6.2f} us/call | one call between events {e0.elapsed_time(e1)*1e3:6.2f} us", flush=True)
    f.close()
