"""GPU: the `Tahoe MODEL DATA` clone (include/BaseTahoeTest.h over the C ABI) runs the reference's
flow on text files and prints the reference's lines; every strategy that runs matches the CPU results."""
import os
import re
import subprocess

import pytest

from tahoe_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _exe():
    exe = os.path.join(ROOT, "tahoe_b200", "bin", "Tahoe")
    if not os.path.exists(exe):
        from tahoe_b200 import build
        build.build()
    return exe


def _check(out, expect_strategies):
    assert "Loading model..." in out and "Predict on CPU to get standard results..." in out
    assert "Results are incorrect" not in out
    ran = re.findall(r"Exec.Time/Sample on strategy (\d) is ([0-9.]+) us", out)
    assert {int(s) for s, _ in ran} >= set(expect_strategies)
    assert out.count("Results are correct") == len(ran) + 1          # + the baseline line
    assert re.search(r"Performance model choose #[1-5] strategy\.", out)
    assert re.search(r"Tahoe brings [0-9.]+x speedup\.", out)


def test_cli_on_golden_text_files():
    g = os.path.join(ROOT, "tests", "golden")
    r = subprocess.run([_exe(), os.path.join(g, "tiny_model.txt"), os.path.join(g, "tiny_data.txt")],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    _check(r.stdout, {1, 2, 3, 4, 5})


def test_cli_on_generated_files(tmp_path):
    T, d, C, R = 60, 7, 18, 3000
    nodes = synth.gen_dense_forest(31, T, d, C, leaf_prob=0.05)
    data = synth.gen_data(32, R, C, miss_frac=0.01)
    m, dfile = str(tmp_path / "m.txt"), str(tmp_path / "d.txt")
    synth.write_model_text(m, nodes, T, d)
    synth.write_data_text(dfile, data, -999.0)
    r = subprocess.run([_exe(), m, dfile], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    _check(r.stdout, {1, 2, 3, 4, 5})
    bad = subprocess.run([_exe(), m], capture_output=True, text=True)
    assert bad.returncode == 2 and "Usage" in bad.stdout
