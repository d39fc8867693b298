"""Fused global-norm clip + Adam over flat buffers (K6).

Drop-in for the optimizer the reference builds at ``agents/impala/builder.py:43-44``
(``torch.optim.Adam(params, lr=1e-4, eps=1e-5)``) together with the ``clip_grad_norm_`` that precedes every
``optimizer.step()`` (``agents/impala/learning.py:172-176``, ``agents/ppo/learning.py:136-140``): parameters,
gradients and both Adam moments live in flat fp32 buffers and one kernel launch does norm -> clip -> update.
"""
from typing import Iterable

import torch

from . import ops
from .parallel import flat_view


class FusedClipAdam(torch.optim.Optimizer):
    """``torch.optim.Adam`` constructor surface (``lr, betas, eps``; no weight decay / amsgrad, as the reference).

    On construction the parameters are re-pointed at views of ONE flat buffer (values preserved).  The
    learners pass the flat gradient bucket (``parallel.GradBucket``) to ``step_flat``; a plain ``step()``
    gathers ``p.grad`` first and is there for API compatibility."""

    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999),
                 eps: float = 1e-8, max_grad_norm: float = 0.0):
        # torch.optim.Adam's full set of group keys (fixed at its defaults: no weight decay / amsgrad, as the reference),
        # so that a checkpoint written here loads into torch's Adam and the other way round
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False, maximize=False,
                                      foreach=None, capturable=False, differentiable=False, fused=None,
                                      decoupled_weight_decay=False))
        if len(self.param_groups) != 1:
            raise RuntimeError("FusedClipAdam: one parameter group only")
        self.max_grad_norm = float(max_grad_norm)
        self.params = [p for p in self.param_groups[0]["params"] if p.requires_grad]
        if not self.params or not all(p.is_cuda and p.dtype == torch.float32 for p in self.params):
            raise RuntimeError("FusedClipAdam needs float32 CUDA parameters (no CPU fallback)")
        dev = self.params[0].device
        n = sum(p.numel() for p in self.params)
        self.flat = torch.empty(n, dtype=torch.float32, device=dev)
        off = 0
        for p in self.params:
            v = flat_view(self.flat, off, p)          # keeps a channels_last weight channels_last
            v.copy_(p.data)
            p.data = v
            off += p.numel()
        self.exp_avg = torch.zeros_like(self.flat)
        self.exp_avg_sq = torch.zeros_like(self.flat)
        self.total_norm = torch.zeros(1, dtype=torch.float32, device=dev)
        self.step_count = 0
        self._own_grads = None
        # graph-capturable mode (``capturable_()``): Adam's step and K8's flag generation live in device memory
        self.device_counters = None

    def capturable_(self, bucket=None):
        """Switch to device-resident counters so that ``step_flat`` / ``step_peer`` can be captured in a CUDA graph and
        replayed: a step number passed as a launch argument would be frozen into the graph.  Call it outside capture, on
        every rank at the same point; ``bucket`` (a ``PeerGradBucket``) hands over its flag generation."""
        if self.device_counters is None:
            gen = getattr(bucket, "_generation", 0) if bucket is not None else 0
            self.device_counters = torch.tensor([self.step_count, gen], dtype=torch.int64, device=self.flat.device)
        return self

    def sync_step_count(self) -> int:
        """Host copy of the device-resident step (synchronises); keeps ``state_dict()`` truthful in capturable mode."""
        if self.device_counters is not None:
            self.step_count = int(self.device_counters[0].item())
        return self.step_count

    def step_flat(self, flat_grads: torch.Tensor, max_grad_norm=None) -> torch.Tensor:
        """One clip + Adam step from a flat gradient buffer laid out like ``self.params``.
        ``flat_grads`` is clipped in place; returns the pre-clip global norm (1-element tensor)."""
        g = self.param_groups[0]
        self.step_count += 1
        mgn = self.max_grad_norm if max_grad_norm is None else float(max_grad_norm)
        return ops.clip_adam_step(self.flat, flat_grads, self.exp_avg, self.exp_avg_sq, self.step_count,
                                  g["lr"], g["betas"][0], g["betas"][1], g["eps"], mgn, self.total_norm,
                                  device_counters=self.device_counters)

    def step_peer(self, bucket, max_grad_norm=None) -> torch.Tensor:
        """K8: allreduce (over the peer-mapped buckets of ``parallel.PeerGradBucket``) + clip + Adam in ONE launch, no
        NCCL call.  The reduced, clipped gradient lands in ``bucket.reduced``, the summed scalar tail in
        ``bucket.extra_reduced``; returns the pre-clip global norm.  Every rank must call it once per step."""
        g = self.param_groups[0]
        self.step_count += 1
        mgn = self.max_grad_norm if max_grad_norm is None else float(max_grad_norm)
        return ops.peer_allreduce_clip_adam_step(self.flat, bucket.bucket_ptrs, bucket.flag_ptrs, bucket.rank, bucket.reduced,
                                                 bucket.extra_reduced, self.exp_avg, self.exp_avg_sq, self.step_count,
                                                 g["lr"], g["betas"][0], g["betas"][1], g["eps"], mgn, self.total_norm,
                                                 bucket.reduced_ptrs, bucket.cta_flag_ptrs, bucket.mc_bucket_ptr,
                                                 bucket.mc_reduced_ptr, flag_generation=bucket.next_generation(),
                                                 device_counters=self.device_counters)

    # -- checkpoint / resume: the layout torch.optim.Adam uses, so either optimizer loads the other's state ---------------
    def _moment_views(self):
        off, out = 0, []
        for p in self.params:
            out.append((p, flat_view(self.exp_avg, off, p), flat_view(self.exp_avg_sq, off, p)))
            off += p.numel()
        return out

    def state_dict(self):
        """``torch.optim.Adam``'s format: per-parameter ``step`` / ``exp_avg`` / ``exp_avg_sq`` (copies of the flat moment
        buffers) -- the base class would otherwise save an EMPTY state and a resumed run would restart the moments and the
        bias correction."""
        self.sync_step_count()
        for p, m, v in self._moment_views():
            self.state[p] = {"step": torch.tensor(float(self.step_count)), "exp_avg": m.detach().clone(),
                             "exp_avg_sq": v.detach().clone()}
        try:
            return super().state_dict()
        finally:
            self.state.clear()

    def load_state_dict(self, state_dict):
        """Accepts a ``FusedClipAdam`` or a ``torch.optim.Adam`` state dict over the same parameters."""
        super().load_state_dict(state_dict)
        steps = set()
        with torch.no_grad():
            for p, m, v in self._moment_views():
                st = self.state.get(p)
                if not st:
                    continue
                m.copy_(st["exp_avg"])
                v.copy_(st["exp_avg_sq"])
                steps.add(int(float(st["step"])))
        if len(steps) > 1:
            raise RuntimeError("FusedClipAdam.load_state_dict: parameters carry different step counts %s" % sorted(steps))
        if steps:
            self.step_count = steps.pop()
            if self.device_counters is not None:
                self.device_counters[0] = self.step_count
        self.state.clear()

    @torch.no_grad()
    def step(self, closure=None):
        if closure is not None:
            raise RuntimeError("FusedClipAdam: closures are not supported")
        if self._own_grads is None:
            self._own_grads = torch.zeros_like(self.flat)
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is not None and p.grad.data_ptr() != self._own_grads[off:off + n].data_ptr():
                # through the parameter's own strides: a channels_last weight sits in the flat buffers in its physical
                # order, so a logical-order flatten of the gradient would pair the wrong elements
                flat_view(self._own_grads, off, p).copy_(p.grad)
            elif p.grad is None:
                # a zero gradient (the moments still decay), where torch.optim.Adam would skip the parameter; the
                # reference's learners produce every gradient on every step
                self._own_grads[off:off + n].zero_()
            off += n
        self.step_flat(self._own_grads)
        return None
