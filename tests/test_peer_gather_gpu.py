"""GPU, two devices: the ranks' score slices are gathered on rank 0 through CUDA IPC peer copies
(tahoe_b200/shard.py PeerGather) -- no NCCL anywhere: the process group is gloo and carries only the
forest broadcast and the pickled IPC handle.  Skipped on a one-GPU box."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
T, D, C, R = 64, 8, 18, 200_003


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    import torch
    import torch.distributed as dist
    from tahoe_b200 import api, shard, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nodes = synth.gen_dense_forest(5, T, D, C, leaf_prob=0.05) if rank == 0 else None
        sf = shard.ShardedForest(nodes, T, D, C, missing=-999.0)
        pg = shard.PeerGather(R)
        x = torch.from_numpy(synth.gen_data(6, pg.hi - pg.lo, C, miss_frac=0.01, row0=pg.lo)).cuda()
        for s in (api.SPLIT_SHARED_FOREST, api.DIRECT):
            sf.predict_local(x, pg.local, strategy=s)
            pg.push()
            torch.cuda.synchronize()
            dist.barrier()
            if rank == 0:
                np.save(out_path + f".{s}.npy", pg.full.cpu().numpy())
            dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_peer_gather_two_gpus(oracle, tmp_path):
    import torch
    import torch.multiprocessing as mp
    from tahoe_b200 import api, synth
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    out = str(tmp_path / "full")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    nodes = synth.gen_dense_forest(5, T, D, C, leaf_prob=0.05)
    want = oracle.predict_dense(nodes, T, D, synth.gen_data(6, R, C, miss_frac=0.01), -999.0)
    for s in (api.SPLIT_SHARED_FOREST, api.DIRECT):
        got = np.load(out + f".{s}.npy")
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))     # thread = sample strategies: the CPU's bits
