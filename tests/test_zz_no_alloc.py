"""tahoe_predict_no_alloc (include/tahoe_b200.h): the shared-memory-forest predict on the library's static
device scratch, for callers that cannot call cudaMalloc.  No reference counterpart.
Named test_zz_* so that pytest collects it LAST.

Through the C ABI, against the existing entry point (tahoe_predict, SPLIT_SHARED_FOREST): same bits --
thread = sample strategies sum in tree order -- on two calls in a row that share the static scratch (the
second smaller and with other data, so stale rank tiles / flag words of the first would show), then the
sparse (quantised tops) forest within the re-ordered-sum tolerance, and the size limit."""
import numpy as np
import pytest

from tahoe_b200 import api, synth
from tests.helpers import bits_equal

pytestmark = pytest.mark.gpu
SCRATCH = 16 << 20      # TAHOE_NO_ALLOC_SCRATCH_BYTES


def _max_rows(cols):
    t = SCRATCH // (cols * 128 + 4)
    while (t * 4 + 255) // 256 * 256 + t * cols * 128 > SCRATCH:
        t -= 1
    return t * 32


def test_no_alloc_matches_predict_twice_in_a_row():
    import torch
    T, d, C = 96, 9, 18
    nodes = synth.gen_dense_forest(901, T, d, C, leaf_prob=0.03)
    f = api.Forest.from_dense(nodes, T, d, C, missing=-999.0)
    assert f.applicable(api.SPLIT_SHARED_FOREST, 1000)
    calls = [(150_003, 0.01, 902), (40_001, 0.0, 903), (_max_rows(C), 0.001, 904)]
    for rows, miss, seed in calls:
        x = api.gen_data_device(seed, rows, C, miss_frac=miss)
        want = f.predict(x, strategy=api.SPLIT_SHARED_FOREST)
        got = torch.full((rows,), float("nan"), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]    # device-wide: a cudaMalloc inside the call would show
        f.predict_no_alloc(x, got)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before
        assert bits_equal(got.cpu().numpy(), want.cpu().numpy()), rows
        assert f.launches() >= 2                      # k_quantize + k_smem_forest ran
    # one row over the limit: a clear error, nothing launched
    rows = _max_rows(C) + 1
    x = torch.zeros((rows, C), dtype=torch.float32, device="cuda")
    out = torch.zeros(rows, dtype=torch.float32, device="cuda")
    with pytest.raises(api.TahoeError) as e:
        f.predict_no_alloc(x, out)
    assert e.value.status == api.ERR_INVALID and str(_max_rows(C)) in str(e.value)


def test_no_alloc_sparse_forest_and_not_applicable(oracle):
    import torch
    sp, offs, w, _ = synth.gen_sparse_forest(911, 64, 14, 50)
    f = api.Forest.from_sparse(sp, offs, 50, weights=w, missing=-999.0)
    assert f.info().sparse_top_active == 1
    for rows, seed in ((30_011, 912), (2_049, 913)):
        x = api.gen_data_device(seed, rows, 50, miss_frac=0.005)
        want = f.predict(x, strategy=api.SPLIT_SHARED_FOREST).cpu().numpy()
        got = torch.empty(rows, dtype=torch.float32, device="cuda")
        f.predict_no_alloc(x, got)
        torch.cuda.synchronize()
        # same kernel, same schedule: reproducible bit for bit (the deep walks' sum order is fixed per launch shape)
        assert bits_equal(got.cpu().numpy(), want)
    # a forest without the rank-quantised layout (100 columns > 64): a status, not a crash
    wide = synth.gen_dense_forest(921, 8, 4, 100)
    fw = api.Forest.from_dense(wide, 8, 4, 100, missing=-999.0)
    x = torch.zeros((64, 100), dtype=torch.float32, device="cuda")
    with pytest.raises(api.NotApplicable):
        fw.predict_no_alloc(x, torch.zeros(64, dtype=torch.float32, device="cuda"))
