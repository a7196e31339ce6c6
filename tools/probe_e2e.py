"""Where the end-to-end step goes on config 2: H2D alone, device predict alone, tahoe_predict_host."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tahoe_b200 import api, synth
cfg = "C2"; c = synth.CONFIGS[cfg]; rows = int(sys.argv[1]) if len(sys.argv) > 1 else c["rows"]
nodes = synth.gen_dense_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
f = api.Forest.from_dense(nodes, c["trees"], c["depth"], c["cols"], missing=-999.0)
d = api.gen_data_device(synth.DATA_SEED[cfg], rows, c["cols"])
h = torch.empty((rows, c["cols"]), dtype=torch.float32, pin_memory=True); h.copy_(d.cpu())
hp = torch.empty(rows, dtype=torch.float32, pin_memory=True)
p = torch.empty(rows, dtype=torch.float32, device="cuda")
def timed(fn, n=5):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
t_h2d = timed(lambda: d.copy_(h, non_blocking=True))
print(f"H2D {h.numel()*4/1e6:.0f} MB: {t_h2d:.3f} ms = {h.numel()*4/t_h2d/1e6:.1f} GB/s")
t_d2h = timed(lambda: hp.copy_(p, non_blocking=True))
print(f"D2H {hp.numel()*4/1e6:.0f} MB: {t_d2h:.3f} ms")
print(f"device predict: {timed(lambda: f.predict(d, p, strategy=api.SPLIT_SHARED_FOREST)):.3f} ms")
hn, hpn = h.numpy(), hp.numpy()
print(f"predict_host: {timed(lambda: f.predict_host(hn, hpn, strategy=api.SPLIT_SHARED_FOREST)):.3f} ms")
