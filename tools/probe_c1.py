"""config 1 (SHARED_DATA on 50K x 3072): kernel time under the plan the environment asks for
(TAHOE_SD_TEAM / TAHOE_SD_BUFS / TAHOE_SD_TOP_LEVELS): python tools/probe_c1.py [rows]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tahoe_b200 import api, synth
c = synth.CONFIGS["C1"]; rows = int(sys.argv[1]) if len(sys.argv) > 1 else c["rows"]
trees = int(os.environ.get("PROBE_TREES", c["trees"])); depth = int(os.environ.get("PROBE_DEPTH", c["depth"]))   # a lighter walk: the kernel's streaming ceiling
nodes = synth.gen_dense_forest(synth.FOREST_SEED["C1"], trees, depth, c["cols"])
f = api.Forest.from_dense(nodes, trees, depth, c["cols"], missing=-999.0)
inf = f.info()
d = api.gen_data_device(synth.DATA_SEED["C1"], rows, c["cols"])
p = torch.empty(rows, dtype=torch.float32, device="cuda")
ref = f.predict(d, strategy=api.DIRECT)
for _ in range(3): f.predict(d, p, strategy=api.SHARED_DATA)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best, tot, n = 1e9, 0.0, 20
for _ in range(n):
    e0.record(); f.predict(d, p, strategy=api.SHARED_DATA); e1.record(); torch.cuda.synchronize()
    t = e0.elapsed_time(e1); best = min(best, t); tot += t
err = float((p.double() - ref.double()).abs().max())
env = {k: os.environ.get(k) for k in ("TAHOE_SD_TEAM", "TAHOE_SD_BUFS", "TAHOE_SD_TOP_LEVELS", "TAHOE_SD_NO_L2_AHEAD") if os.environ.get(k)}
print(f"C1 rows={rows} trees={trees} depth={depth} {env} teams={inf.sd_teams} team_warps={inf.sd_team_warps} top={inf.sd_top_levels} blocked={inf.sd_blocked} "
      f"mean={tot/n:.4f} ms best={best:.4f} ms hbm_only={(rows*12292+306600)/(tot/n*1e-3)/1e9/6552.3:.3f} max|sd-direct|={err:.2e}", flush=True)
