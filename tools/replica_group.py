"""One process per GPU, no data-path collective: the reference scales by running more copies of the
container behind a round-robin router (APIs/Charts/templates/routing.yml:20-28,
deployment.yml:13-17), so the only cross-rank traffic here is measurement plumbing - a barrier
on both sides of a timed region and the max of the per-rank device times.

`ReplicaGroup` wraps torch.distributed for that (NCCL on the GPU box, gloo in the CPU tests); with
a world of one it never touches torch.distributed.
"""
from __future__ import annotations

import os
from typing import Optional, Sequence


class ReplicaGroup:
    def __init__(self, backend: Optional[str] = None, device_index: Optional[int] = None):
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.backend = backend
        self._dist = None
        self._device = "cpu"
        if backend == "nccl":  # GPU timing: barriers also drain the device, with one rank too
            import torch
            idx = self.local_rank if device_index is None else device_index
            self._device = torch.device("cuda", idx)
        if self.world > 1:
            import torch
            import torch.distributed as dist
            if backend is None:
                backend = "gloo"
            self.backend = backend
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            kw = {"device_id": self._device} if backend == "nccl" else {}
            dist.init_process_group(backend, rank=self.rank, world_size=self.world, **kw)
            self._dist = dist

    # ------------------------------------------------------------------ plumbing
    def _sync_device(self):
        if self._device != "cpu":
            import torch
            torch.cuda.synchronize(self._device)

    def barrier(self):
        """device idle on every rank before anyone proceeds (sync, barrier, sync)."""
        self._sync_device()
        if self._dist is not None:
            self._dist.barrier()
        self._sync_device()

    def max_over_ranks(self, x: float) -> float:
        if self._dist is None:
            return float(x)
        import torch
        t = torch.tensor([float(x)], dtype=torch.float64, device=self._device)
        self._dist.all_reduce(t, op=self._dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, x: float) -> float:
        if self._dist is None:
            return float(x)
        import torch
        t = torch.tensor([float(x)], dtype=torch.float64, device=self._device)
        self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM)
        return float(t.item())

    def aggregate_rate(self, units_this_rank: float, elapsed_ms_this_rank: float) -> float:
        """whole-job throughput: the units all ranks processed / the slowest rank's time."""
        total = self.sum_over_ranks(units_this_rank)
        ms = self.max_over_ranks(elapsed_ms_this_rank)
        return total / (ms / 1e3)

    def close(self):
        if self._dist is not None:
            self._dist.destroy_process_group()
            self._dist = None


def shard_range(n_items: int, world: int, rank: int) -> range:
    """contiguous share of `n_items` independent requests for `rank` (sizes differ by at most one)."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} of {world}")
    base, extra = divmod(max(n_items, 0), world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def merge_shards(parts: Sequence[Sequence]) -> list:
    """inverse of shard_range: rank-ordered shares back into request order."""
    out = []
    for p in parts:
        out.extend(p)
    return out
