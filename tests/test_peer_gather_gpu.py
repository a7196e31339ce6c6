"""GPU, two devices: the ranks' final scores are stored by the kernels' own epilogue straight into
rank 0's buffer, mapped through CUDA IPC (tahoe_predict_to, tahoe_b200/shard.py PeerGather) -- no copy
and no NCCL anywhere: the process group is gloo and carries only the forest broadcast and the pickled
IPC handle.  The gathered buffer must equal the single-GPU result bit for bit.  With two GPUs the ranks sit on
different devices (stores cross NVLink); on a one-GPU box both ranks share cuda:0 -- two processes, the same IPC
mapping and epilogue-store path, no link in between."""
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
    torch.cuda.set_device(rank % torch.cuda.device_count())
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nodes = synth.gen_dense_forest(5, T, D, C, leaf_prob=0.05) if rank == 0 else None
        sf = shard.ShardedForest(nodes, T, D, C, missing=-999.0)
        pg = shard.PeerGather(R)
        x = torch.from_numpy(synth.gen_data(6, pg.hi - pg.lo, C, miss_frac=0.01, row0=pg.lo)).cuda()
        for s in (api.SPLIT_SHARED_FOREST, api.DIRECT, api.SHARED_DATA):
            if rank == 0:
                pg.full.fill_(float("nan"))            # every row must be (re)written by some rank's kernel
                torch.cuda.synchronize()
            dist.barrier()
            pg.predict(sf.forest, x, strategy=s)
            torch.cuda.synchronize()
            dist.barrier()
            if rank == 0:
                np.save(out_path + f".{s}.npy", pg.full.cpu().numpy())
            dist.barrier()
        if rank != 0:
            pg.close()
        dist.barrier()
        pg.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_peer_gather_two_gpus(oracle, tmp_path):
    import torch
    import torch.multiprocessing as mp
    from tahoe_b200 import api, synth
    if torch.cuda.device_count() < 1:
        pytest.skip("needs a GPU")
    out = str(tmp_path / "full")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    nodes = synth.gen_dense_forest(5, T, D, C, leaf_prob=0.05)
    want = oracle.predict_dense(nodes, T, D, synth.gen_data(6, R, C, miss_frac=0.01), -999.0)
    # the same forest on ONE GPU, all rows: the gathered buffer must hold exactly these bits
    f = api.Forest.from_dense(nodes, T, D, C, missing=-999.0)
    x = torch.from_numpy(synth.gen_data(6, R, C, miss_frac=0.01)).cuda()
    for s in (api.SPLIT_SHARED_FOREST, api.DIRECT, api.SHARED_DATA):
        got = np.load(out + f".{s}.npy")
        single = f.predict(x, strategy=s).cpu().numpy()
        assert np.array_equal(got.view(np.uint32), single.view(np.uint32)), api.STRATEGY_NAMES[s]
        if s != api.SHARED_DATA:   # thread = sample strategies: the CPU's bits
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
