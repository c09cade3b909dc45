"""Multi-rank plumbing for bench.py and the tests: one process per GPU (torchrun), utterances
sharded by rank, no collective on the data path. torch.distributed is used only
for the barrier and for the max / sum over ranks of the timing and sample counters; `nccl` on the
GPU box, `gloo` in the CPU tests."""
from __future__ import annotations

import contextlib
import os
import sys


@contextlib.contextmanager
def _stdout_to_stderr():
    """NCCL prints its version banner on stdout when the first communicator is created; bench.py's
    stdout must carry exactly one JSON line, so the banner is sent to stderr (file-descriptor level:
    the print comes from C)."""
    sys.stdout.flush()
    saved = os.dup(1)
    try:
        os.dup2(2, 1)
        yield
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)


def dist_env():
    """(rank, world, local_rank) from the torchrun environment; (0, 1, 0) when not under torchrun."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def rank_utterance_range(utts_per_rank: int, rank: int):
    """Weak scaling: rank r owns utterances [r * n, (r + 1) * n) of the global synthetic batch."""
    return rank * utts_per_rank, (rank + 1) * utts_per_rank


class Reducer:
    """barrier / max / sum over ranks. backend None = single process (no torch import)."""

    def __init__(self, world: int, backend: str | None, local_rank: int = 0):
        self.world, self.backend, self.dist, self.device = world, backend, None, "cpu"
        if world > 1:
            import torch
            import torch.distributed as dist
            if backend == "nccl":
                torch.cuda.set_device(local_rank)
                self.device = f"cuda:{local_rank}"
                with _stdout_to_stderr():
                    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
                    dist.barrier()  # creates the communicator now, inside the redirect
            else:
                dist.init_process_group(backend)
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def _reduce(self, x: float, op_name: str) -> float:
        if self.dist is None:
            return float(x)
        import torch
        t = torch.tensor([x], dtype=torch.float64, device=self.device)
        self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op_name))
        return float(t.item())

    def max(self, x: float) -> float:
        return self._reduce(x, "MAX")

    def sum(self, x: float) -> float:
        return self._reduce(x, "SUM")

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()
            self.dist = None
