"""SPLIT_SHARED_FOREST on a synthetic dense forest of any shape: python tools/probe_deep_dense.py TREES DEPTH COLS ROWS
(TAHOE_PAIR / TAHOE_SMEM_PLAN choose the kernel and the plan)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tahoe_b200 import api, synth
T, D, C, R = (int(x) for x in sys.argv[1:5])
nodes = synth.gen_dense_forest(777 + D, T, D, C)
f = api.Forest.from_dense(nodes, T, D, C, missing=-999.0)
inf = f.info()
d = api.gen_data_device(4242, R, C)
p = torch.empty(R, dtype=torch.float32, device="cuda")
ref = f.predict(d, strategy=api.DIRECT)
for _ in range(2): f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST)
torch.cuda.synchronize(); f.set_profiling(True)
for _ in range(3): f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST)
q, t, n = f.profile_read()
print(f"PAIR={os.environ.get('TAHOE_PAIR','auto')} {T}x d{D} x{C} rows={R} plan={os.environ.get('TAHOE_SMEM_PLAN','auto')} -> warps={inf.smem_warps} staged={inf.smem_staged} "
      f"G={inf.smem_group} tpc={inf.trees_per_chunk} chunks={inf.num_chunks} quantize={q/n:.3f} ms traverse={t/n:.3f} ms max|split-direct|={float((p.double()-ref.double()).abs().max()):.2e}")
