"""Per-kernel timing of the sparse shared-memory path on config 4: python tools/probe_sparse.py [rows]
(TAHOE_SPARSE_CUT / TAHOE_SPARSE_TPC are read when the forest is packed)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tahoe_b200 import api, synth
cfg = "C4"; c = synth.CONFIGS[cfg]; rows = int(sys.argv[1]) if len(sys.argv) > 1 else c["rows"]
sp, offs, w, lv = synth.gen_sparse_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
t0 = time.time()
f = api.Forest.from_sparse(sp, offs, c["cols"], weights=w, missing=-999.0)
pack_s = time.time() - t0
inf = f.info()
d = api.gen_data_device(synth.DATA_SEED[cfg], rows, c["cols"])
p = torch.empty(rows, dtype=torch.float32, device="cuda")
ref = f.predict(d, strategy=api.DIRECT)
for _ in range(2): f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST)
torch.cuda.synchronize(); f.set_profiling(True)
for _ in range(4): f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST)
q, t, n = f.profile_read()
err = float((p.double() - ref.double()).abs().max())
print(f"C4 rows={rows} cut={inf.sparse_cut} tpc={inf.trees_per_chunk} chunks={inf.num_chunks} deep={inf.sparse_deep_walks:.4f} "
      f"pack={pack_s:.1f}s tiles={q/n:.3f} ms traverse={t/n:.3f} ms max|split-direct|={err:.2e}", flush=True)
