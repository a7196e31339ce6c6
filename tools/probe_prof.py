"""Per-kernel timing of the shared-memory-forest path: python tools/probe_prof.py [C2|C3] [rows]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tahoe_b200 import api, synth
cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
c = synth.CONFIGS[cfg]; rows = int(sys.argv[2]) if len(sys.argv) > 2 else c["rows"]
nodes = synth.gen_dense_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
f = api.Forest.from_dense(nodes, c["trees"], c["depth"], c["cols"], missing=-999.0)
inf = f.info()
d = api.gen_data_device(synth.DATA_SEED[cfg], rows, c["cols"])
p = torch.empty(rows, dtype=torch.float32, device="cuda")
for _ in range(3): f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST)
torch.cuda.synchronize(); f.set_profiling(True)
for _ in range(5): f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST)
q, t, n = f.profile_read()
print(f"TL={os.environ.get('TAHOE_TOP_LEVELS','auto')} S={os.environ.get('TAHOE_CLUSTER','auto')} {cfg} rows={rows} plan={os.environ.get('TAHOE_SMEM_PLAN','auto')} -> warps={inf.smem_warps} staged={inf.smem_staged} G={inf.smem_group} tpc={inf.trees_per_chunk} chunks={inf.num_chunks} col_shift={inf.col_shift} quantize={q/n:.3f} ms traverse={t/n:.3f} ms  checksum={float(p.double().sum()):.6f}")
