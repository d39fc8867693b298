"""Multi-GPU plumbing: one process per GPU, rollout batches sharded along B.

The reference has no multi-GPU path.  V-trace / GAE and every per-(b,t) loss term are
independent per batch column (the reference itself writes that as ``vmap`` over dim 0,
``agents/impala/learning.py:18``), so the loss kernels run rank-local with ``n_global_b`` and
the only exchange is ONE allreduce per step: the flattened model gradient with the 8 loss/metric
scalars riding in its tail (NCCL over NVLink/NVSwitch on GPUs, gloo in the CPU tests).
"""
import os
import warnings
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


def flat_view(flat: torch.Tensor, off: int, p: torch.Tensor) -> torch.Tensor:
    """View of ``flat[off : off + p.numel()]`` with ``p``'s shape AND memory format: a channels_last convolution weight
    stays channels_last when its storage moves into a flat bucket (physical ``[O][3][3][I]``, what K10/K11 read and
    write), so no layout copy is needed per step; everything else is a plain contiguous view.  Adam and the allreduce
    are elementwise, so the order of elements inside the slice is free as long as parameter, gradient and moments
    use the same one -- they do, all three are built by this function from the same strides."""
    chunk = flat[off:off + p.numel()]
    if p.dim() == 4 and not p.is_contiguous() and p.is_contiguous(memory_format=torch.channels_last):
        return chunk.as_strided(p.shape, p.stride())
    return chunk.view_as(p)


class GradBucket:
    """All parameter gradients as views into one flat fp32 buffer + ``n_extra`` tail floats, so a step
    needs exactly one allreduce (630,352 params = 2.52 MB for the IMPALA-CNN: latency-bound, one bucket)."""

    def __init__(self, params: Iterable[torch.nn.Parameter], n_extra: int = 8, reserve: int = 0, store=None):
        self.params = [p for p in params if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        dev = self.params[0].device
        # reserve / store: see PeerGradBucket (extra floats behind the bucket; a pre-allocated symmetric buffer)
        self.store = torch.zeros(n + n_extra + reserve, dtype=torch.float32, device=dev) if store is None else store
        self.flat = self.store[:n + n_extra]
        self.n, self.n_extra = n, n_extra
        self.views, off = [], 0
        for p in self.params:
            self.views.append(flat_view(self.flat, off, p))
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


class PeerGradBucket(GradBucket):
    """A ``GradBucket`` whose buffer is mapped into every rank's address space (NVLink / NVSwitch peer access), for the
    fused allreduce + clip + Adam kernel (K8, ``ops.peer_allreduce_clip_adam_step``): no NCCL call on the step.
    Transport (``B200RL_PEER_TRANSPORT`` = ``auto`` | ``symm`` | ``ipc``): torch symmetric memory when available -- it
    also yields the NVLS multicast address the kernel's ``multimem`` route needs -- else CUDA IPC handles exchanged
    through ``all_gather_object`` (unicast peer loads / stores only).  Buffer layout: ``n`` gradients, ``n_extra`` scalars, padding to 16 bytes, 16 flag words, 8,192 per-CTA
    flag words and the ``n``-float reduced-gradient buffer the ranks write into each other (two-shot route)."""

    FLAG_WORDS = 16
    CTA_FLAG_WORDS = 1024 * 8          # two-shot route: per-CTA flags, [kMaxOptCtas][8]

    def __init__(self, params: Iterable[torch.nn.Parameter], n_extra: int = 8, group=None):
        from . import ops
        plist = [p for p in params if p.requires_grad]
        n = sum(p.numel() for p in plist)
        pad = (-(n + n_extra)) % 4
        n_pad = n + ((-n) % 4)
        reserve = pad + self.FLAG_WORDS + self.CTA_FLAG_WORDS + n_pad
        if not plist[0].is_cuda:
            raise RuntimeError("PeerGradBucket needs CUDA parameters (peer memory is a GPU feature)")
        self.group = group
        self.rank, self.world_size = world(group)
        if self.world_size > 8:
            raise RuntimeError("PeerGradBucket: at most 8 ranks (one NVSwitch box)")
        transport = os.environ.get("B200RL_PEER_TRANSPORT", "auto")
        store, self._symm, self.mc_ptr = None, None, 0
        if self.world_size > 1 and transport != "ipc":
            try:
                import torch.distributed._symmetric_memory as symm_mem
                store = symm_mem.empty(n + n_extra + reserve, dtype=torch.float32, device=plist[0].device)
                store.zero_()
                self._symm = symm_mem.rendezvous(store, group if group is not None else dist.group.WORLD)
                self.mc_ptr = int(self._symm.multicast_ptr or 0)
            except Exception as e:   # noqa: BLE001 -- any failure here is symmetric across ranks (same build, same box)
                if transport == "symm":
                    raise RuntimeError("symmetric-memory transport unavailable: %s" % e)
                store, self._symm, self.mc_ptr = None, None, 0
        super().__init__(plist, n_extra, reserve=reserve, store=store)
        self.transport = "symm" if self._symm is not None else "ipc"
        self.flag_offset_bytes = (n + n_extra + pad) * 4
        self._imports = []
        if self._symm is not None:
            self.bucket_ptrs = [int(p) for p in self._symm.buffer_ptrs]
            if self.bucket_ptrs[self.rank] != self.store.data_ptr():
                raise RuntimeError("symmetric memory: local buffer pointer mismatch")
        else:
            mine = ops.ipc_export(self.store)
            gathered = [None] * self.world_size
            if self.world_size > 1:
                dist.all_gather_object(gathered, mine, group=group)
            else:
                gathered[0] = mine
            self.bucket_ptrs = []
            for r, (handle, offset) in enumerate(gathered):
                if r == self.rank:
                    self.bucket_ptrs.append(self.store.data_ptr())
                else:
                    ptr = ops.ipc_import(handle, offset)
                    self._imports.append((ptr, offset))
                    self.bucket_ptrs.append(ptr)
        self.flag_ptrs = [p + self.flag_offset_bytes for p in self.bucket_ptrs]
        cta_flag_off = self.flag_offset_bytes + self.FLAG_WORDS * 4
        reduced_off = cta_flag_off + self.CTA_FLAG_WORDS * 4
        self.cta_flag_ptrs = [p + cta_flag_off for p in self.bucket_ptrs]
        self.reduced_ptrs = [p + reduced_off for p in self.bucket_ptrs]
        self.mc_bucket_ptr = self.mc_ptr                                    # NVLS multicast views of the same offsets
        self.mc_reduced_ptr = self.mc_ptr + reduced_off if self.mc_ptr else 0
        self.reduced = self.store[reduced_off // 4: reduced_off // 4 + self.n]   # reduced, clipped gradient (peer-written in two-shot)
        self.extra_reduced = torch.zeros(n_extra, dtype=torch.float32, device=self.store.device)
        if self.world_size > 1:
            dist.barrier(group=group)   # every rank has mapped every bucket before the first kernel touches one

    def next_generation(self) -> int:
        """The value the cross-rank flags of this bucket carry on the next K8 launch: a monotonic counter owned by the
        bucket (its flag areas were zeroed when it was created, under a barrier), advanced by one per launch on every rank --
        NOT the optimizer's step, which a resumed or recreated optimizer may set to anything."""
        self._generation = getattr(self, "_generation", 0) + 1
        return self._generation

    def allreduce(self, group=None) -> None:
        raise RuntimeError("PeerGradBucket: the allreduce is fused into FusedClipAdam.step_peer (K8)")

    def close(self) -> None:
        from . import ops
        for ptr, offset in self._imports:
            ops.ipc_close(ptr, offset)
        self._imports = []


def make_bucket(params, group=None, fused_optimizer: bool = False, n_extra: int = 8) -> GradBucket:
    """The learners' gradient bucket: peer-mapped (K8) when the optimizer is ``FusedClipAdam``, more than one rank runs
    on CUDA and ``B200RL_PEER_ALLREDUCE`` is not ``0``; otherwise the plain bucket (one NCCL / gloo allreduce per step).
    ``B200RL_PEER_ALLREDUCE=1`` makes a failed IPC mapping an error instead of a warning + NCCL."""
    plist = [p for p in params if p.requires_grad]
    mode = os.environ.get("B200RL_PEER_ALLREDUCE", "auto")
    _, ws = world(group)
    if fused_optimizer and ws > 1 and plist[0].is_cuda and mode != "0":
        try:
            return PeerGradBucket(plist, n_extra, group)
        except RuntimeError as e:   # export / import failures are symmetric across ranks (same allocator, same box)
            if mode == "1":
                raise
            warnings.warn("peer-mapped gradient bucket unavailable (%s); using the NCCL allreduce" % e)
    return GradBucket(plist, n_extra)
