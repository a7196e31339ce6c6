"""One config-4 predict through the sparse shared-memory path, for ncu: python tools/run_qtop.py ROWS REPS"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tahoe_b200 import api, synth
cfg = "C4"; c = synth.CONFIGS[cfg]; rows = int(sys.argv[1]); reps = int(sys.argv[2])
sp, offs, w, lv = synth.gen_sparse_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
f = api.Forest.from_sparse(sp, offs, c["cols"], weights=w, missing=-999.0)
d = api.gen_data_device(synth.DATA_SEED[cfg], rows, c["cols"])
p = torch.empty(rows, dtype=torch.float32, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(reps):
    e0.record(); f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST); e1.record(); torch.cuda.synchronize()
    print(f"C4 rows={rows} rep {i}: {e0.elapsed_time(e1):.3f} ms", flush=True)
