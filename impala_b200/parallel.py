"""Multi-GPU plumbing: one process per GPU, rollout batches sharded along B.

The reference has no multi-GPU path.  V-trace / GAE and every per-(b,t) loss term are
independent per batch column (the reference itself writes that as ``vmap`` over dim 0,
``agents/impala/learning.py:18``), so the loss kernels run rank-local with ``n_global_b`` and
the only exchange is ONE allreduce per step: the flattened model gradient with the 8 loss/metric
scalars riding in its tail (NCCL over NVLink/NVSwitch on GPUs, gloo in the CPU tests).
"""
from typing import Iterable, Optional

import torch
import torch.distributed as dist


def world(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_columns(B_global: int, rank: int, world_size: int):
    """Columns ``[lo, hi)`` of rank ``rank`` (contiguous, remainder to the low ranks)."""
    base, rem = divmod(B_global, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class GradBucket:
    """All parameter gradients as views into one flat fp32 buffer + ``n_extra`` tail floats, so a step
    needs exactly one allreduce (630,352 params = 2.52 MB for the IMPALA-CNN: latency-bound, one bucket)."""

    def __init__(self, params: Iterable[torch.nn.Parameter], n_extra: int = 8):
        self.params = [p for p in params if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        dev = self.params[0].device
        self.flat = torch.zeros(n + n_extra, dtype=torch.float32, device=dev)
        self.n, self.n_extra = n, n_extra
        self.views, off = [], 0
        for p in self.params:
            self.views.append(self.flat[off:off + p.numel()].view_as(p))
            off += p.numel()

    @property
    def grads(self) -> torch.Tensor:
        return self.flat[:self.n]

    @property
    def extra(self) -> torch.Tensor:
        return self.flat[self.n:]

    def attach(self) -> None:
        """Zero the bucket and point every ``p.grad`` at its slice (replaces ``zero_grad``)."""
        self.flat.zero_()
        for p, v in zip(self.params, self.views):
            p.grad = v

    def allreduce(self, group=None) -> None:
        """Sum over ranks (gradients were scaled for the global mean by ``n_global_b``)."""
        if world(group)[1] > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)

    def clip_(self, max_norm: float) -> torch.Tensor:
        """``clip_grad_norm_`` on the reduced bucket (norm after reduction = single-GPU semantics,
        reference ``agents/impala/learning.py:172-173``); no host sync."""
        g = self.grads
        total = torch.linalg.vector_norm(g)
        g.mul_(torch.clamp(max_norm / (total + 1e-6), max=1.0))
        return total
