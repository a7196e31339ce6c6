"""tahoe_predict_host against tahoe_predict on the same rows: python tools/probe_host_vs_device.py CFG rows"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tahoe_b200 import api, synth
cfg = sys.argv[1]; c = synth.CONFIGS[cfg]; rows = int(sys.argv[2])
nodes = synth.gen_dense_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
f = api.Forest.from_dense(nodes, c["trees"], c["depth"], c["cols"], missing=-999.0)
d = api.gen_data_device(synth.DATA_SEED[cfg], rows, c["cols"])
dev = f.predict(d, strategy=api.SPLIT_SHARED_FOREST).cpu().numpy()
ref = f.predict(d, strategy=api.DIRECT).cpu().numpy()
h = torch.empty((rows, c["cols"]), dtype=torch.float32, pin_memory=True); h.copy_(d)
hp = torch.empty(rows, dtype=torch.float32, pin_memory=True)
for it in range(int(sys.argv[3]) if len(sys.argv) > 3 else 3):
    f.predict_host(h, hp, strategy=api.SPLIT_SHARED_FOREST)
    got = hp.numpy().copy()
    bad = np.nonzero(got.view(np.uint32) != dev.view(np.uint32))[0]
    print(f"{cfg} rows={rows} tiles_env={os.environ.get('TAHOE_QUANT_TILES')} slab_env={os.environ.get('TAHOE_HOST_SLAB')} iter {it}: device==direct {np.array_equal(dev.view(np.uint32), ref.view(np.uint32))} "
          f"host!=device rows {bad.size}" + (f" first {bad[:4]} last {bad[-4:]} tiles {np.unique(bad // 32).size} max|diff| {np.abs(got[bad]-dev[bad]).max():.3g}" if bad.size else ""), flush=True)
