"""Ablation of the I8 layout's probability reorder (Struct.h:1775-1825 rule) and simhash tree order on the
strategy that uses them (SHARED_DATA, lane = tree): python tools/probe_reorder.py T DEPTH COLS ROWS
Run once per setting of TAHOE_NO_REORDER / TAHOE_NO_SIMHASH (read when the forest is packed).

The synthetic forests draw their cover weights at random; for the reorder to mean anything the weights here are
CALIBRATED to the data: features are U(-1,1), so a node with threshold t sends a row right with probability
(1 - t) / 2, and a node's weight is the product of those along its path (what a trainer's cover would be)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tahoe_b200 import api, synth
T, d, C, R = (int(x) for x in sys.argv[1:5])
nodes = synth.gen_dense_forest(31337, T, d, C)
npt = (1 << (d + 1)) - 1
w = nodes["weight"].reshape(T, npt); thr = nodes["val"].reshape(T, npt)
w[:, 0] = 1.0
for k in range((1 << d) - 1):
    pr = np.clip((1.0 - thr[:, k]) / 2.0, 0.0, 1.0)
    w[:, 2 * k + 1] = w[:, k] * (1.0 - pr)
    w[:, 2 * k + 2] = w[:, k] * pr
f = api.Forest.from_dense(nodes, T, d, C, missing=-999.0)
x = api.gen_data_device(4711, R, C)
p = torch.empty(R, dtype=torch.float32, device="cuda")
ref = f.predict(x, strategy=api.DIRECT)
for _ in range(3): f.predict(x, p, strategy=api.SHARED_DATA)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for _ in range(10):
    e0.record(); f.predict(x, p, strategy=api.SHARED_DATA); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
inf = f.info()
print(f"{T}x d{d} x{C} rows={R} reorder={'off' if os.environ.get('TAHOE_NO_REORDER') else 'on'} "
      f"simhash={'off' if os.environ.get('TAHOE_NO_SIMHASH') else 'on'} exchanged={inf.exchanged_nodes} "
      f"top={inf.sd_top_levels} blocked={inf.sd_blocked} SHARED_DATA best={best:.4f} ms "
      f"max|sd-direct|={float((p.double()-ref.double()).abs().max()):.2e}", flush=True)
