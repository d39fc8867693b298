"""Drop-in for the reference's ``ImpalaLearner`` (``agents/impala/learning.py:86-177``)."""
import time
from typing import Dict, Optional

import torch

from .. import ops, optim, parallel
from ._graph import GraphedStepMixin


def make_discount(done: torch.Tensor, gamma: float = 0.99) -> torch.Tensor:
    """``discount_t = logical_not(d) * gamma`` (reference ``ImpalaActor._make_replay``, ``:79``)."""
    return torch.logical_not(done) * gamma


class ImpalaLearner(GraphedStepMixin):
    """Same constructor, ``prepare()``, ``train_step()`` and ``_train_step(batch)`` as the reference.

    Differences, all on the hot path: the replay buffer is device-resident (``impala_b200.replay``),
    the loss + its backward w.r.t. the model outputs is one fused kernel, metrics stay on the device
    (no per-metric sync), and with ``torch.distributed`` initialised the batch is a rank-local shard of
    a global batch of ``batch_size * world_size`` rollouts with one gradient allreduce per step."""

    def __init__(self, model, replay_buffer, optimizer: torch.optim.Optimizer, batch_size: int = 8,
                 max_grad_norm: float = 0.5, entropy_coeff: float = 0.01,
                 learning_starts: Optional[int] = None, model_push_period: int = 4,
                 lambda_: float = 1.0, process_group=None) -> None:
        self._model = model
        self._replay_buffer = replay_buffer
        self._optimizer = optimizer
        self._batch_size = batch_size
        self._max_grad_norm = max_grad_norm
        self._entropy_coeff = entropy_coeff
        self._learning_starts = learning_starts
        self._model_push_period = model_push_period
        self._lambda = lambda_
        self._group = process_group
        self._device = None
        self._step_count = 0
        self._bucket = None

    def device(self) -> torch.device:
        if self._device is None:
            self._device = next(self._model.parameters()).device
        return self._device

    def prepare(self):
        self._replay_buffer.warm_up(self._learning_starts)

    def train_step(self):
        if self._graph is not None:                 # capture_graph(): ONE launch per step (agents/_graph.py)
            t0 = time.perf_counter()
            metrics, dt = self._replay_step()
            update_time = self._push_model()
            # same keys as the eager path; sampling and the gradient step are one launch here, so they share its host time
            n_samples = self._batch_size * self._batch_shape[1]
            metrics["debug/replay_sample_per_second"] = metrics["debug/gradient_per_second"] = n_samples / max(dt * 1000, 1e-9)
            metrics["debug/total_time"] = (time.perf_counter() - t0) * 1000
            metrics["debug/forward_dt"] = dt * 1000 / self._batch_size
            metrics["debug/update_time"] = update_time
            return metrics
        t0 = time.perf_counter()
        _, batch, _ = self._replay_buffer.sample(self._batch_size)
        n_samples = self._batch_size * batch[0].shape[1]
        t1 = time.perf_counter()
        metrics = self._train_step(batch)
        t2 = time.perf_counter()
        update_time = self._push_model()
        metrics["debug/replay_sample_per_second"] = n_samples / max((t1 - t0) * 1000, 1e-9)
        metrics["debug/gradient_per_second"] = n_samples / max((t2 - t1) * 1000, 1e-9)
        metrics["debug/total_time"] = (time.perf_counter() - t0) * 1000
        metrics["debug/forward_dt"] = (t2 - t1) * 1000 / self._batch_size
        metrics["debug/update_time"] = update_time
        return metrics

    @staticmethod
    def _collate(batch):
        """Stacked fields ``[B,T,...]`` (device replay) or the reference's list of B items (``:142``)."""
        if isinstance(batch[0], (list, tuple)):
            batch = [torch.stack(x) for x in zip(*batch)]
        return [x.squeeze(-1) if (x.dim() == 3 and x.shape[-1] == 1) else x for x in batch]

    def _train_step(self, batch) -> Dict[str, torch.Tensor]:
        if self._bucket is None:
            self._bucket = parallel.make_bucket(self._model.parameters(), self._group,
                                                isinstance(self._optimizer, optim.FusedClipAdam))
            if isinstance(self._optimizer, optim.FusedClipAdam) and \
                    [id(q) for q in self._bucket.params] != [id(q) for q in self._optimizer.params]:
                raise RuntimeError("FusedClipAdam must be built from model.parameters() (same order as the gradient bucket)")
        bucket = self._bucket
        mark = self._mark
        mark("start")
        bucket.attach()                                                   # == zero_grad (:141)
        s, a, r, discount_t, pi_ref = self._collate(batch)
        dev = self.device()
        s, a, r, discount_t, pi_ref = (x.to(dev, non_blocking=True) for x in (s, a, r, discount_t, pi_ref))
        B, T = s.shape[0], s.shape[1]
        _, world_size = parallel.world(self._group)
        mark("collate+h2d")
        pi, values = self._model.forward(s.flatten(0, 1))                 # :143
        pi = pi.reshape(B, T, -1)
        values = values.reshape(B, T)
        mark("model fwd")
        scalars, dlogits, dvalues = ops.impala_loss_fwdbwd(               # :146-160 fused
            pi, values, a, r, discount_t, pi_ref, self._entropy_coeff, self._lambda,
            n_global_b=B * world_size)
        mark("K1 loss fwd+bwd")
        torch.autograd.backward([pi, values], [dlogits, dvalues])
        mark("model bwd")
        bucket.extra.copy_(scalars)
        if isinstance(bucket, parallel.PeerGradBucket):                   # allreduce + :172-176 in one launch (K8)
            mark("grad allreduce")
            grad_norm = self._optimizer.step_peer(bucket, self._max_grad_norm)[0]
            sc = bucket.extra_reduced
        else:
            bucket.allreduce(self._group)                                 # grads + scalars, one collective
            mark("grad allreduce")
            if isinstance(self._optimizer, optim.FusedClipAdam):           # :172-176 in one launch (K6)
                grad_norm = self._optimizer.step_flat(bucket.grads, self._max_grad_norm)[0]
            else:
                grad_norm = bucket.clip_(self._max_grad_norm)                 # :172-173
                self._optimizer.step()                                        # :176
            sc = bucket.extra
        mark("clip+adam")
        return {
            "train/loss": sc[ops.S_LOSS], "train/entropy": sc[ops.S_ENTROPY], "train/td": sc[ops.S_TD],
            "train/pg": sc[ops.S_PG], "train/kl": sc[ops.S_KL], "train/ratio": sc[ops.S_RATIO],
            "train/grad_norm": grad_norm,
        }

    # -- optional per-phase device timing (the reference only has host perf_counter deltas, :133-137) --------
    profile = False

    def _mark(self, name: str) -> None:
        if not self.profile:
            return
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        if name == "start":
            self._events = []
        self._events.append((name, ev))

    def breakdown_ms(self) -> Dict[str, float]:
        """Device time of each phase of the last profiled ``_train_step`` (synchronises)."""
        torch.cuda.synchronize()
        ev = getattr(self, "_events", [])
        return {n1: e0.elapsed_time(e1) for (_, e0), (n1, e1) in zip(ev, ev[1:])}
