"""Sample-sharded multi-GPU plumbing: one process per GPU, the forest replicated
once, rows split into contiguous ranges, no collective on the predict path.

The reference is single-GPU (BaseTahoeTest.h:78, 645-647); this is the one parallel strategy the
path admits: pred[r] depends only on row r and the read-only forest (BaseTahoeTest.h:462-466).
torch.distributed is plumbing only: a one-time broadcast of the AoS forest (NCCL on GPUs, gloo in the
CPU tests) and an optional, untimed gather of the score slices.
"""
from __future__ import annotations

import numpy as np

from .synth import DENSE_NODE


def row_range(rank: int, world: int, total_rows: int) -> tuple[int, int]:
    """Contiguous range [g*R/G, (g+1)*R/G) of rank g (64-bit arithmetic; ranges tile [0, R) exactly)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    lo = (int(total_rows) * rank) // world
    hi = (int(total_rows) * (rank + 1)) // world
    return lo, hi


def broadcast_forest(nodes: np.ndarray | None, num_nodes: int, src: int = 0, device=None) -> np.ndarray:
    """One-time broadcast of the dense AoS forest (12 B/node) from `src` to every rank."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        assert nodes is not None
        return np.ascontiguousarray(nodes, dtype=DENSE_NODE)
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    buf = torch.empty(num_nodes * DENSE_NODE.itemsize, dtype=torch.uint8, device=dev)
    if dist.get_rank() == src:
        src_np = np.ascontiguousarray(nodes, dtype=DENSE_NODE).view(np.uint8).reshape(-1)
        assert src_np.shape[0] == buf.shape[0], "node count mismatch"
        buf.copy_(torch.from_numpy(src_np))
    dist.broadcast(buf, src=src)
    return buf.cpu().numpy().view(DENSE_NODE).copy()


def gather_scores(local, total_rows: int, dst: int = 0):
    """Untimed convenience: concatenate the ranks' score slices (in rank order = row order) on `dst`.
    `local` is a torch tensor on the backend's device; returns the full tensor on dst, None elsewhere."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = [row_range(g, world, total_rows) for g in range(world)]
    assert local.shape[0] == sizes[rank][1] - sizes[rank][0], "slice does not match this rank's row range"
    longest = max(hi - lo for lo, hi in sizes)
    padded = torch.zeros(longest, dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    parts = [torch.empty_like(padded) for _ in range(world)] if rank == dst else None
    dist.gather(padded, parts, dst=dst)
    if rank != dst:
        return None
    return torch.cat([parts[g][: hi - lo] for g, (lo, hi) in enumerate(sizes)])


class ShardedForest:
    """One handle per process/GPU over the same forest; predict() runs the local rows only."""

    def __init__(self, nodes, num_trees: int, depth: int, num_cols: int, *, src: int = 0, **kw):
        from . import api
        n = num_trees * ((1 << (depth + 1)) - 1)
        self.nodes = broadcast_forest(nodes, n, src=src)
        self.forest = api.Forest.from_dense(self.nodes, num_trees, depth, num_cols, **kw)

    def predict_local(self, data, preds=None, **kw):
        return self.forest.predict(data, preds, **kw)


class PeerGather:
    """Gather of the ranks' score slices WITHOUT a collective and WITHOUT a copy: `dst`
    owns one device buffer for the whole batch (tahoe_gather_buffer_create), every other rank maps it through
    CUDA IPC once, with its own device current (tahoe_gather_buffer_open; the 64-byte handle travels as a
    pickled object over the process group -- gloo is enough), and
    `tahoe_predict_to` makes the kernels' last-chunk epilogue store each final score straight into
    that buffer over NVLink / NVSwitch (4 bytes per row, posted writes, never read back).  The partial
    sums the multi-pass strategies carry between forest chunks stay in a local scratch.  The only
    synchronisation is the barrier the caller already has.

        pg = PeerGather(total_rows)                  # once
        pg.predict(forest, x_local)                  # final scores land in dst's buffer, at this rank's rows
        torch.cuda.synchronize(); dist.barrier()     # dst may now read pg.full
    """

    def __init__(self, total_rows: int, dst: int = 0):
        import torch
        import torch.distributed as dist
        from . import api
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.dst, self.total = dst, int(total_rows)
        self.lo, self.hi = row_range(self.rank, self.world, self.total)
        dev = torch.device("cuda", torch.cuda.current_device())
        self.owner = self.rank == dst
        box = [None]
        if self.owner:      # a cudaMalloc of its own, so that the IPC handle names exactly this buffer
            self.ptr, box[0] = api.gather_buffer_create(self.total)
        if self.world > 1:
            dist.broadcast_object_list(box, src=dst)
            if not self.owner:
                # dst's memory, mapped for the kernels of THIS rank's GPU (opened with this device current)
                self.ptr = api.gather_buffer_open(box[0])
        # dst reads the gathered scores through a tensor view of its own buffer; the other ranks only ever hand the
        # raw pointer (+ their row offset) to tahoe_predict_to as d_out
        self.full = torch.as_tensor(api.DevicePointer(self.ptr, self.total), device=dev) if self.owner else None
        self.out = self.ptr + 4 * self.lo              # this rank's rows of the gathered buffer
        self.partial = torch.empty(self.hi - self.lo, dtype=torch.float32, device=dev)   # local scratch

    def predict(self, forest, data, **kw):
        """One predict of this rank's rows; the final store of every row goes to dst's buffer."""
        return forest.predict_to(data, self.out, partial=self.partial, **kw)

    def close(self):
        """Collective in spirit: every rank closes its mapping before the owner frees the buffer (callers barrier)."""
        from . import api
        if getattr(self, "ptr", None):
            self.full = None
            api.gather_buffer_release(self.ptr, self.owner)
            self.ptr = None
