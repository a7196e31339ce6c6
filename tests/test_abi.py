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


# ---- fast text loaders + binary cache ------------------------------------------------
TRICKY = ["0.5", "-0.25", "  1.5e-3", "+7", "1e", "1e+", "12abc", "abc", "", ".", "-.5", "5.", "0x1p-3", "inf", "-inf",
          "nan", "1e400", "1e-400", "123456789012345678901234567890", "0.1", "0.30000001192092896", "3.4028235e38",
          "1.17549435e-38", "1.401298464324817e-45", "7.0064923216240854e-46", "-0.0", "0e999", "1.0000000596046448",
          "1.00000005960464477539062500000000001", "8388608.5", "16777217", "  \t 42.5xyz", "4.5\r", "\r", "1,5",
          "9007199254740993", "0.000000000000000000000000000001", "1e22", "1e23", "123456789.123456789e-5", "- 5"]


def _write(path, lines, trailing_newline=True):
    with open(path, "wb") as fh:
        fh.write("\n".join(lines).encode() + (b"\n" if trailing_newline else b""))


@pytest.mark.parametrize("threads,min_range", [(1, 1 << 16), (7, 16), (64, 1)])
def test_fast_loaders_match_the_sequential_reference_loop(tmp_path, monkeypatch, threads, min_range):
    """The mapped multi-threaded loaders against the one-thread fgets/atof loop (the reference's own
    algorithm, kept as TAHOE_LOADER_SEQUENTIAL=1) on files full of edge cases: garbage, blank lines,
    hex floats, inf/nan, double-rounding traps, CRLF, no trailing newline, short files."""
    rng = np.random.default_rng(5)
    vals = [repr(float(np.float32(x))) for x in rng.uniform(-1, 1, 400)] + \
           ["%.17g" % x for x in rng.normal(0, 1e3, 200)] + ["%e" % x for x in rng.normal(0, 1e-20, 100)] + TRICKY
    rng.shuffle(vals)
    R, Cc = 37, 20
    data_lines = [str(R), str(Cc), "-999.0"] + [vals[i % len(vals)] for i in range(R * Cc)]
    T, d = 3, 3
    model_lines = [str(T), str(d + 1)]
    for i in range(T * 15):
        model_lines += [str(int(rng.integers(0, 20))) if i % 7 else "  7junk", vals[i % len(vals)],
                        str(int(rng.integers(0, 2))), vals[(3 * i + 1) % len(vals)], str(int(i % 15 >= 7))]
    cases = {
        "full": (data_lines, model_lines, True),
        "no_newline": (data_lines, model_lines, False),
        "short": (data_lines[: 3 + R * Cc // 2], model_lines[: 2 + 5 * 20 + 3], True),
        "short_no_newline": (data_lines[: 3 + 11], model_lines[: 2 + 4], False),
        "header_only": (data_lines[:2], model_lines[:1], True),
    }
    for name, (dl, ml, nl) in cases.items():
        dp, mp = str(tmp_path / f"{name}_data.txt"), str(tmp_path / f"{name}_model.txt")
        _write(dp, dl, nl)
        _write(mp, ml, nl)
        monkeypatch.setenv("TAHOE_LOADER_SEQUENTIAL", "1")
        want_d, want_m = api.load_data_text(dp), api.load_model_text(mp)
        monkeypatch.delenv("TAHOE_LOADER_SEQUENTIAL")
        monkeypatch.setenv("TAHOE_LOADER_THREADS", str(threads))
        monkeypatch.setenv("TAHOE_LOADER_MIN_RANGE", str(min_range))
        got_d, got_m = api.load_data_text(dp), api.load_model_text(mp)
        assert got_d[0].shape == want_d[0].shape and got_d[0].tobytes() == want_d[0].tobytes(), name
        assert np.float32(got_d[1]).tobytes() == np.float32(want_d[1]).tobytes(), name
        assert got_m[1:] == want_m[1:] and got_m[0].tobytes() == want_m[0].tobytes(), name


def test_fast_loaders_fall_back_on_lines_fgets_would_split(tmp_path, monkeypatch):
    """A line of 1024+ bytes is read by the reference's fgets in pieces (each piece one value): the fast
    path detects it and the sequential loop takes over."""
    lines = ["2", "3", "-999.0", "1.5", " " * 1500 + "2.5", "3.5", "4.5", "5.5", "6.5"]
    dp = str(tmp_path / "long.txt")
    _write(dp, lines)
    monkeypatch.setenv("TAHOE_LOADER_SEQUENTIAL", "1")
    want = api.load_data_text(dp)
    monkeypatch.delenv("TAHOE_LOADER_SEQUENTIAL")
    got = api.load_data_text(dp)
    assert got[0].tobytes() == want[0].tobytes()
    assert want[0].ravel()[1] == 0.0 and want[0].ravel()[2] == 2.5        # the long line counted twice


def test_fast_loaders_match_the_reference_goldens(golden_dir):
    z = np.load(os.path.join(golden_dir, "text.npz"))
    for stem in ("tiny", "quirks"):
        nodes, t, d = api.load_model_text(os.path.join(golden_dir, f"{stem}_model.txt"))
        data, missing = api.load_data_text(os.path.join(golden_dir, f"{stem}_data.txt"))
        assert (t, d) == (int(z[f"{stem}_trees"]), int(z[f"{stem}_depth"]))
        assert nodes.tobytes() == z[f"{stem}_nodes"].tobytes() and data.tobytes() == z[f"{stem}_data"].tobytes()
        assert np.float32(missing) == z[f"{stem}_missing"]


def test_binary_cache_round_trip_and_corruption(tmp_path, golden_dir):
    nodes, t, d = api.load_model_text(os.path.join(golden_dir, "tiny_model.txt"))
    data, missing = api.load_data_text(os.path.join(golden_dir, "tiny_data.txt"))
    mp, dp = str(tmp_path / "m.tb2"), str(tmp_path / "d.tb2")
    api.save_model_binary(mp, nodes, t, d)
    api.save_data_binary(dp, data, missing)
    n2, t2, d2 = api.load_model_binary(mp)
    x2, m2 = api.load_data_binary(dp)
    assert (t2, d2) == (t, d) and n2.tobytes() == nodes.tobytes()
    assert x2.shape == data.shape and x2.tobytes() == data.tobytes() and m2 == missing
    with pytest.raises(api.TahoeError):                       # wrong kind
        api.load_model_binary(dp)
    raw = bytearray(open(dp, "rb").read())
    raw[100] ^= 0x40                                          # one flipped payload bit: checksum
    open(dp, "wb").write(raw)
    with pytest.raises(api.TahoeError):
        api.load_data_binary(dp)
    open(mp, "wb").write(open(mp, "rb").read()[:-5])          # truncated
    with pytest.raises(api.TahoeError):
        api.load_model_binary(mp)
    with pytest.raises(api.TahoeError):
        api.load_model_binary(str(tmp_path / "missing.tb2"))
    # empty shapes survive
    api.save_data_binary(dp, np.zeros((0, 4), np.float32), -1.0)
    x3, m3 = api.load_data_binary(dp)
    assert x3.shape == (0, 4) and m3 == -1.0


@pytest.mark.skipif(not os.path.exists("/root/reference/Struct.h"), reason="needs the reference tree (this container only)")
def test_integration_stub_compiles_against_the_reference_headers(tmp_path):
    """the binding a reference maintainer would add (`b200_forest : forest`
    over the C ABI) is extracted from the document and compiled + linked against the reference's own
    Struct.h, in place, with the reference's build flags -- the drop-in claim, checked."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    stub = re.search(r"```cpp\n(// tahoe_b200_forest\.h.*?)```", doc, re.S).group(1)
    (tmp_path / "tahoe_b200_forest.h").write_text(stub)
    (tmp_path / "main.cu").write_text(
        '#include "tahoe_b200_forest.h"\n'
        "int main() { forest_params_t ps{}; b200_forest* f = new b200_forest; (void)ps; delete f; return 0; }\n")
    from tahoe_b200 import build
    lib = os.path.dirname(build.build())
    exe = str(tmp_path / "stub_main")
    cmd = ["nvcc", "-include", "cfloat", "-include", "algorithm", "-I/root/reference/inc", "-I/root/reference",
           "-I" + os.path.join(root, "include"), "-std=c++17", "-w", "-gencode", "arch=compute_100a,code=sm_100a",
           str(tmp_path / "main.cu"), "-o", exe, "-L" + lib, "-ltahoe_b200", "-Xlinker", "-rpath", "-Xlinker", lib]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert os.path.exists(exe)
