"""C-ABI surface (no GPU): the library loads, exports every symbol include/tahoe_b200.h declares,
its text loaders agree with the reference's, the host mirror compiles, and the product never reaches
into oracle/."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

from tahoe_b200 import api, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "tahoe_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tahoe_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = api.lib()
    names = _declared_symbols()
    assert len(names) >= 18 and "tahoe_predict" in names and "tahoe_forest_create_dense" in names
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in include/tahoe_b200.h but not exported"
    assert lib.tahoe_version().decode().startswith("tahoe_b200")


def test_pod_layouts_match_the_reference():
    assert synth.DENSE_NODE.itemsize == 12 and synth.SPARSE_NODE.itemsize == 12      # Struct.h:44-54
    assert ctypes.sizeof(api.ForestParams) == 40                                      # Struct.h:166-189


def test_text_loaders_match_reference_goldens(golden_dir):
    z = np.load(os.path.join(golden_dir, "text.npz"))
    nodes, T, d = api.load_model_text(os.path.join(golden_dir, "tiny_model.txt"))
    assert (T, d) == (int(z["tiny_trees"]), int(z["tiny_depth"]))
    assert nodes.tobytes() == z["tiny_nodes"].tobytes()
    data, missing = api.load_data_text(os.path.join(golden_dir, "tiny_data.txt"))
    assert data.tobytes() == z["tiny_data"].tobytes() and missing == float(z["tiny_missing"])
    qn, qT, qd = api.load_model_text(os.path.join(golden_dir, "quirks_model.txt"))
    assert qn.tobytes() == z["quirks_nodes"].tobytes()
    qdat, qm = api.load_data_text(os.path.join(golden_dir, "quirks_data.txt"))
    assert qdat.tobytes() == z["quirks_data"].tobytes()
    with pytest.raises(api.TahoeError):
        api.load_model_text(os.path.join(golden_dir, "does_not_exist.txt"))


def test_strategy_ids_follow_the_reference_dispatch():
    # selected_algorithm 0..4, Struct.h:2168-2179
    assert (api.DIRECT, api.SHARED_DATA_PER_THREAD, api.SHARED_DATA, api.SHARED_FOREST,
            api.SPLIT_SHARED_FOREST) == (0, 1, 2, 3, 4)


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "tahoe_b200")):
        if any(part in dirpath for part in ("/build", "/lib", "/bin", "__pycache__")):
            continue
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, fn), errors="replace").read()
                assert not re.search(r"(import|include|dlopen|CDLL)[^\n]*oracle", text), f"{fn} reaches into oracle/"
    hdr = open(os.path.join(ROOT, "include", "BaseTahoeTest.h")).read()
    assert not re.search(r"(include|dlopen)[^\n]*oracle", hdr)


def test_host_mirror_compiles_and_cli_fails_loudly_without_gpu(tmp_path):
    """include/BaseTahoeTest.h + the CLI clone build against the C ABI; with no GPU the CLI says so
    and exits non-zero (no CPU fallback)."""
    import torch
    exe = str(tmp_path / "Tahoe")
    cmd = ["nvcc", "-O1", "-std=c++17", "-Wno-deprecated-gpu-targets", "-o", exe,
           os.path.join(ROOT, "tahoe_b200", "cli", "main.cpp"),
           "-L" + os.path.join(ROOT, "tahoe_b200", "lib"), "-ltahoe_b200",
           "-Xlinker", "-rpath", "-Xlinker", os.path.join(ROOT, "tahoe_b200", "lib")]
    subprocess.run(cmd, check=True, capture_output=True)
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by test_cli_gpu")
    g = os.path.join(ROOT, "tests", "golden")
    r = subprocess.run([exe, os.path.join(g, "tiny_model.txt"), os.path.join(g, "tiny_data.txt")],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "FAIL" in r.stdout
