"""Per-kernel timing of the sparse quantised-tops path (QT) on config 4: python tools/probe_qtop.py [rows]
(TAHOE_QTOP_DEPTH / TAHOE_SMEM_PLAN="warps,staged,group" are read when the forest is packed; without a GPU
only the plan is printed)"""
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
plan = (f"D={inf.sparse_top_depth} active={inf.sparse_top_active} tpc={inf.sparse_top_trees_per_chunk} "
        f"chunks={inf.sparse_top_chunks} deep={inf.sparse_top_deep_walks:.4f} rank_bits={inf.rank_bits} "
        f"col_shift={inf.col_shift} warps={inf.smem_warps} staged={inf.smem_staged} G={inf.smem_group} pack={pack_s:.1f}s")
if not torch.cuda.is_available():
    print("C4 plan:", plan); sys.exit(0)
d = api.gen_data_device(synth.DATA_SEED[cfg], rows, c["cols"])
p = torch.empty(rows, dtype=torch.float32, device="cuda")
ref = f.predict(d, strategy=api.DIRECT)
for _ in range(2): f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST)
torch.cuda.synchronize(); f.set_profiling(True)
for _ in range(4): f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST)
q, t, n = f.profile_read()
err = float((p.double() - ref.double()).abs().max())
print(f"C4 rows={rows} {plan} quantize={q/n:.3f} ms traverse={t/n:.3f} ms max|split-direct|={err:.2e}", flush=True)
