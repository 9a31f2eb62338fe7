"""Instance sharding for the multi-GPU crowd.

Instances are independent units: no joint, clip or vertex of one instance depends on another, so the path shards
by instance range with replicated assets and NO data-path collective.  torch.distributed (NCCL on the GPU box,
gloo in the CPU tests) is used only for the barrier around the timed region and the max-over-ranks reduction.
"""
from __future__ import annotations

import os


def shard_range(n_total: int, world: int, rank: int):
    """GPU g of G owns [g*ceil(N/G), min(N, (g+1)*ceil(N/G)))  ->  (first, count)."""
    per = -(-n_total // world)
    first = min(n_total, rank * per)
    return first, max(0, min(n_total, first + per) - first)


def env_rank():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def open_gathered_pool(group: "RankGroup", ctx, skel: int, n_total: int, consumer: int = 0):
    """one palette pool of n_total rows on rank `consumer`'s GPU, opened on every other rank through CUDA
    IPC.  Each rank then attaches its shard's batch at `shard_range(n_total, world, rank)[0]`, so that the pose kernel stores
    its palettes straight into the consumer's memory over NVLink; torch.distributed only carries the 64-byte handle."""
    from . import capi
    if group.rank == consumer:
        pool = ctx.pool(skel, n_total)
        handle = pool.export()
    else:
        pool, handle = None, b"\0" * capi.PCL_POOL_HANDLE_BYTES
    handle = group.broadcast_bytes(handle, capi.PCL_POOL_HANDLE_BYTES, src=consumer)
    if pool is None:
        pool = ctx.pool(skel, n_total, handle)
    return pool


class RankGroup:
    """Thin wrapper: barrier + max/sum over ranks; degenerates to a no-op for a single process."""

    def __init__(self, backend: str = "nccl", device=None):
        self.rank, self.local, self.world = env_rank()
        self.dist = None
        self.device = device
        if self.world > 1:
            import torch.distributed as dist
            kw = {}
            if backend == "nccl":
                import torch
                torch.cuda.set_device(self.local)
                self.device = torch.device("cuda", self.local)
                kw["device_id"] = self.device
            dist.init_process_group(backend, **kw)
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        if self.device is not None and str(self.device).startswith("cuda"):
            import torch
            torch.cuda.synchronize()

    def _reduce(self, x: float, op_name: str) -> float:
        if self.dist is None:
            return float(x)
        import torch
        t = torch.tensor([x], dtype=torch.float64, device=self.device if self.device is not None else "cpu")
        self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op_name))
        return float(t.item())

    def max(self, x: float) -> float:
        return self._reduce(x, "MAX")

    def sum(self, x: float) -> float:
        return self._reduce(x, "SUM")

    def broadcast_bytes(self, payload: bytes, nbytes: int, src: int = 0) -> bytes:
        """`payload` (exactly nbytes long) of rank `src` on every rank - e.g. a palette pool's IPC handle."""
        if self.dist is None:
            return bytes(payload)
        import torch
        dev = self.device if self.device is not None else "cpu"
        t = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        if self.rank == src:
            assert len(payload) == nbytes
            t.copy_(torch.frombuffer(bytearray(payload), dtype=torch.uint8))
        self.dist.broadcast(t, src=src)
        return bytes(t.cpu().numpy().tobytes())

    def close(self):
        if self.dist is not None:
            self.dist.barrier()
            self.dist.destroy_process_group()
            self.dist = None
