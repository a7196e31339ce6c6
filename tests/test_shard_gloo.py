"""world_size-2 gloo test (CPU) of the sample-sharded multi-GPU host logic: row ranges tile the batch,
the forest broadcast is bit-exact, and the rank-ordered gather of per-rank scores equals the
single-process result.  The per-rank "predict" here is the oracle (test infrastructure standing in for
the GPU kernel, which the -m gpu tests cover); what is under test is tahoe_b200/shard.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tahoe_b200 import shard, synth

T, D, C, R = 24, 5, 12, 1001


def test_row_ranges_tile_exactly():
    for world in (1, 2, 3, 8):
        for total in (0, 1, 7, 8, 1001, 10_000_000, 3_000_000_000):
            spans = [shard.row_range(g, world, total) for g in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.row_range(2, 2, 10)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle.oracle import Oracle
        n_nodes = T * synth.tree_num_nodes(D)
        nodes = synth.gen_dense_forest(99, T, D, C, leaf_prob=0.1) if rank == 0 else None
        nodes = shard.broadcast_forest(nodes, n_nodes, src=0)
        want_nodes = synth.gen_dense_forest(99, T, D, C, leaf_prob=0.1)
        assert nodes.tobytes() == want_nodes.tobytes()
        lo, hi = shard.row_range(rank, world, R)
        local = synth.gen_data(7, hi - lo, C, miss_frac=0.02, row0=lo)      # this rank's shard only
        got = Oracle().predict_dense(nodes, T, D, local, -999.0)
        full = shard.gather_scores(torch.from_numpy(got), R, dst=0)
        if rank == 0:
            np.save(out_path, full.numpy())
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_rank_shard_equals_single_process(tmp_path, oracle):
    out = str(tmp_path / "scores.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = np.load(out)
    nodes = synth.gen_dense_forest(99, T, D, C, leaf_prob=0.1)
    data = synth.gen_data(7, R, C, miss_frac=0.02)
    want = oracle.predict_dense(nodes, T, D, data, -999.0)
    assert got.shape == (R,) and np.array_equal(got.view(np.uint32), want.view(np.uint32))
