"""Run one strategy on one config a few times (profiling target): python tools/run_one.py C1 2 [rows] [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tahoe_b200 import api, synth
cfg = sys.argv[1]; s = int(sys.argv[2]); c = synth.CONFIGS[cfg]
rows = int(sys.argv[3]) if len(sys.argv) > 3 and int(sys.argv[3]) > 0 else c["rows"]
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
if c["sparse"]:
    sp, offs, w, lv = synth.gen_sparse_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
    f = api.Forest.from_sparse(sp, offs, c["cols"], weights=w, missing=-999.0)
else:
    nodes = synth.gen_dense_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
    f = api.Forest.from_dense(nodes, c["trees"], c["depth"], c["cols"], missing=-999.0)
d = api.gen_data_device(synth.DATA_SEED[cfg], rows, c["cols"])
p = torch.empty(rows, dtype=torch.float32, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(iters):
    e0.record(); f.predict(d, p, strategy=s); e1.record(); torch.cuda.synchronize()
    print(f"{cfg} {api.STRATEGY_NAMES[s]} rows={rows} iter {i}: {e0.elapsed_time(e1):.3f} ms", flush=True)
