"""Loader throughput: lines/s of the data-text and model-text loaders on a config-3-shaped
file, next to the reference's own fgets + atof loop (BaseTahoeTest.h:267-402, through oracle/_ref when
present -- checker code, used here only as the timed baseline) and to the binary cache.

    python tools/loader_throughput.py [rows=1000000] > profiles/r02_loader_throughput.json

Config 3's data file is 280 M lines (10 M rows x 28); the default here is a tenth of it (28 M lines, one line
per value): the loaders are linear in the line count, and the full file is 2.8 GB of text."""
import json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tahoe_b200 import api, synth

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
c = synth.CONFIGS["C3"]
tmp = tempfile.mkdtemp(prefix="tahoe_loader_")
dpath, mpath = os.path.join(tmp, "d.txt"), os.path.join(tmp, "m.txt")
data = synth.gen_data(synth.DATA_SEED["C3"], rows, c["cols"])
np.savetxt(dpath, data.reshape(-1), fmt="%.9g", header=f"{rows}\n{c['cols']}\n-999", comments="")
T = 100   # 100 of config 3's 1000 trees: 100 x 8191 nodes x 5 lines = 4.1 M lines
nodes = synth.gen_dense_forest(synth.FOREST_SEED["C3"], T, c["depth"], c["cols"])
synth.write_model_text(mpath, nodes, T, c["depth"])
dlines, mlines = rows * c["cols"] + 3, nodes.shape[0] * 5 + 2


def timed(fn, reps=2):
    best = float("inf")
    out = None
    for _ in range(reps):
        t0 = time.perf_counter(); out = fn(); best = min(best, time.perf_counter() - t0)
    return best, out


res = {"host_threads": len(os.sched_getaffinity(0)), "data_file": {"rows": rows, "cols": c["cols"], "lines": dlines,
       "bytes": os.path.getsize(dpath)}, "model_file": {"trees": T, "depth": c["depth"], "lines": mlines,
       "bytes": os.path.getsize(mpath)}}
t, (d_par, miss) = timed(lambda: api.load_data_text(dpath))
res["data_parallel"] = {"seconds": t, "lines_per_s": dlines / t}
os.environ["TAHOE_LOADER_SEQUENTIAL"] = "1"
t, (d_seq, _) = timed(lambda: api.load_data_text(dpath), reps=1)
res["data_sequential_same_library"] = {"seconds": t, "lines_per_s": dlines / t}
del os.environ["TAHOE_LOADER_SEQUENTIAL"]
assert d_par.tobytes() == d_seq.tobytes() == data.tobytes()
t, (m_par, _, _) = timed(lambda: api.load_model_text(mpath))
res["model_parallel"] = {"seconds": t, "lines_per_s": mlines / t}
assert m_par.tobytes() == nodes.tobytes()
try:
    from oracle.oracle import RefShim
    if RefShim.available():
        shim = RefShim()
        t, out = timed(lambda: shim.load_data(dpath), reps=1)
        res["data_reference_loop"] = {"seconds": t, "lines_per_s": dlines / t, "what": "generate_data_from_file, BaseTahoeTest.h:354-388 (oracle/_ref)"}
        assert np.asarray(out[0]).tobytes() == data.tobytes()
        t, out = timed(lambda: shim.load_model(mpath), reps=1)
        res["model_reference_loop"] = {"seconds": t, "lines_per_s": mlines / t, "what": "generate_forest_from_file, BaseTahoeTest.h:267-352 (oracle/_ref)"}
except Exception as e:  # noqa: BLE001
    res["reference_loop_error"] = repr(e)
bpath = os.path.join(tmp, "d.bin")
t, _ = timed(lambda: api.save_data_binary(bpath, data, -999.0), reps=1)
res["binary_cache_save"] = {"seconds": t, "gb_per_s": data.nbytes / t / 1e9}
t, (d_bin, _) = timed(lambda: api.load_data_binary(bpath))
res["binary_cache_load"] = {"seconds": t, "gb_per_s": data.nbytes / t / 1e9, "equivalent_lines_per_s": dlines / t}
assert d_bin.tobytes() == data.tobytes()
if "data_reference_loop" in res:
    res["speedup_parallel_vs_reference_loop"] = res["data_reference_loop"]["seconds"] / res["data_parallel"]["seconds"]
    res["config3_full_file_estimate_s"] = {"lines": 280_000_003,
        "reference_loop": 280_000_003 / res["data_reference_loop"]["lines_per_s"],
        "parallel": 280_000_003 / res["data_parallel"]["lines_per_s"],
        "binary_cache": 1.12e9 / (res["binary_cache_load"]["gb_per_s"] * 1e9)}
for p in (dpath, mpath, bpath):
    os.remove(p)
os.rmdir(tmp)
print(json.dumps(res, indent=1))
