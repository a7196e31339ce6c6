"""python tools/run_sparse.py ROWS [iters] -- config 4 through SPLIT_SHARED_FOREST (profiling target)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tahoe_b200 import api, synth
cfg = "C4"; c = synth.CONFIGS[cfg]; rows = int(sys.argv[1]); iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
sp, offs, w, lv = synth.gen_sparse_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
f = api.Forest.from_sparse(sp, offs, c["cols"], weights=w, missing=-999.0)
d = api.gen_data_device(synth.DATA_SEED[cfg], rows, c["cols"])
p = torch.empty(rows, dtype=torch.float32, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(iters):
    e0.record(); f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST); e1.record(); torch.cuda.synchronize()
    print(f"C4 split rows={rows} iter {i}: {e0.elapsed_time(e1):.3f} ms", flush=True)
