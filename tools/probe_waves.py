"""Fixed cost of one k_smem_forest launch on config 2: traverse time for whole waves of tiles
(148 SMs x 16 warps x 32 rows = 75,776 rows per wave), per-kernel CUDA events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tahoe_b200 import api, synth
cfg = "C2"; c = synth.CONFIGS[cfg]
nodes = synth.gen_dense_forest(synth.FOREST_SEED[cfg], c["trees"], c["depth"], c["cols"])
f = api.Forest.from_dense(nodes, c["trees"], c["depth"], c["cols"], missing=-999.0)
wave = 148 * 16 * 32
d = api.gen_data_device(synth.DATA_SEED[cfg], 66 * wave, c["cols"])
for k in (1, 2, 3, 4, 8, 16, 32, 66):
    x = d[: k * wave]; p = torch.empty(k * wave, dtype=torch.float32, device="cuda")
    for _ in range(3): f.predict(x, p, strategy=api.SPLIT_SHARED_FOREST)
    torch.cuda.synchronize(); f.set_profiling(True)
    for _ in range(10): f.predict(x, p, strategy=api.SPLIT_SHARED_FOREST)
    q, t, n = f.profile_read(); f.set_profiling(False)
    print(f"waves={k:3d} rows={k*wave:8d} quantize={q/n:.4f} ms traverse={t/n:.4f} ms  per wave {t/n/k*1e3:.1f} us", flush=True)
