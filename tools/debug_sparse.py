import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tahoe_b200 import api, synth
rows = int(sys.argv[1]); want_ids = int(sys.argv[2]); trees = int(sys.argv[3]) if len(sys.argv) > 3 else 64
sp, offs, w, lv = synth.gen_sparse_forest(4242, trees, 12, 50)
data = synth.gen_data(4343, rows, 50, miss_frac=0.01, nan_frac=0.005)
f = api.Forest.from_sparse(sp, offs, 50, weights=w, missing=-999.0)
i = f.info()
d = torch.from_numpy(data).cuda()
ids = torch.zeros((rows, trees), dtype=torch.int32, device="cuda") if want_ids else None
try:
    if os.environ.get("DIRECT_FIRST"):
        b = f.predict(d, strategy=api.DIRECT); torch.cuda.synchronize()
        print("direct ok", end=" ")
    a = f.predict(d, strategy=api.SPLIT_SHARED_FOREST, leaf_ids=ids); torch.cuda.synchronize()
    print("split ok", end=" ")
    b = f.predict(d, strategy=api.DIRECT); torch.cuda.synchronize()
    print(f"rows={rows} ids={want_ids} trees={trees} cut={i.sparse_cut} tpc={i.trees_per_chunk} chunks={i.num_chunks}: ok maxdiff={float((a-b).abs().max()):.2e}")
except Exception as ex:
    print(f"rows={rows} ids={want_ids} trees={trees} cut={i.sparse_cut} tpc={i.trees_per_chunk} chunks={i.num_chunks}: FAIL {str(ex).splitlines()[0]}")
