"""Drop-in for the reference's ``PPOLearner`` (``agents/ppo/learning.py:75-143``) and the actor-side
target computation (``PPOActorRemote._make_replay``, ``:65-69``)."""
import time
from typing import Dict, Optional

import torch

from .. import ops, optim, parallel, vtrace
from ._graph import GraphedStepMixin


def make_targets(r, done, v, gamma: float = 0.99, lambda_: float = 0.95, time_major: bool = False):
    """Batched ``_make_replay`` targets: ``lambda_returns(r[:-1], discount[:-1], v[1:], lambda_)`` for
    every column at once (the reference loops per rollout on the CPU, and its committed line has a
    NameError -- ``values`` for ``v``).  ``[B,T]`` inputs (``[T,B]`` if ``time_major``) -> ``T-1`` steps."""
    discount = torch.logical_not(done) * gamma
    if time_major:
        return vtrace.lambda_returns(r[:-1], discount[:-1], v[1:], lambda_, time_major=True)
    return vtrace.lambda_returns(r[:, :-1], discount[:, :-1], v[:, 1:], lambda_)


class PPOLearner(GraphedStepMixin):
    """Same constructor, ``prepare()``, ``train_step()`` and ``_train_step(batch)`` as the reference; ``capture_graph()``
    (``agents/_graph.py``) turns the step into one CUDA graph launch."""

    def __init__(self, model, replay_buffer, optimizer: torch.optim.Optimizer, batch_size: int = 256,
                 max_grad_norm: float = 0.5, entropy_coeff: float = 0.01,
                 learning_starts: Optional[int] = None, model_push_period: int = 8,
                 clip_coeff: float = 0.1, process_group=None) -> None:
        self._model = model
        self._replay_buffer = replay_buffer
        self._optimizer = optimizer
        self._batch_size = batch_size
        self._max_grad_norm = max_grad_norm
        self._entropy_coeff = entropy_coeff
        self._learning_starts = learning_starts
        self._model_push_period = model_push_period
        self._clip_coeff = clip_coeff
        self._group = process_group
        self._device = None
        self._step_count = 0
        self._bucket = None
        self.can_train = False

    @property
    def _step_counter(self) -> int:                 # the reference's name for it (agents/ppo/learning.py:98)
        return self._step_count

    def device(self) -> torch.device:
        if self._device is None:
            self._device = next(self._model.parameters()).device
        return self._device

    def prepare(self):
        if not self.can_train:
            self._replay_buffer.warm_up(self._learning_starts)
        self.can_train = True

    def train_step(self):
        t0 = time.perf_counter()
        if self._graph is not None:                 # capture_graph(): ONE launch per step
            metrics, dt = self._replay_step()
            t1 = t0
            t2 = t0 + dt
        else:
            _, batch, _ = self._replay_buffer.sample(self._batch_size)
            t1 = time.perf_counter()
            metrics = self._train_step(batch)
            t2 = time.perf_counter()
        update_time = self._push_model() * 1000                            # :108-112
        metrics["debug/replay_sample_per_second"] = self._batch_size / max(((t1 - t0) or (t2 - t0)) * 1000, 1e-9)
        metrics["debug/gradient_per_second"] = self._batch_size / max((t2 - t1) * 1000, 1e-9)
        metrics["debug/total_time"] = (time.perf_counter() - t0) * 1000
        metrics["debug/forward_dt"] = (t2 - t1) * 1000
        metrics["debug/update_time"] = update_time
        return metrics

    def _train_step(self, batch) -> Dict[str, torch.Tensor]:
        if self._bucket is None:
            self._bucket = parallel.make_bucket(self._model.parameters(), self._group,
                                                isinstance(self._optimizer, optim.FusedClipAdam))
            if isinstance(self._optimizer, optim.FusedClipAdam) and \
                    [id(q) for q in self._bucket.params] != [id(q) for q in self._optimizer.params]:
                raise RuntimeError("FusedClipAdam must be built from model.parameters() (same order as the gradient bucket)")
        bucket = self._bucket
        bucket.attach()
        dev = self.device()
        s, a, v_target, pi_ref = (x.to(dev, non_blocking=True) for x in batch)
        a, v_target = a.reshape(-1), v_target.reshape(-1)
        _, world_size = parallel.world(self._group)
        pi, v = self._model(s)
        scalars, dlogits, dv = ops.ppo_loss_fwdbwd(pi, v, a, v_target, pi_ref, self._entropy_coeff,
                                                   self._clip_coeff, n_global=pi.shape[0] * world_size)
        torch.autograd.backward([pi, v], [dlogits, dv.view_as(v)])        # :134
        bucket.extra.copy_(scalars)
        if isinstance(bucket, parallel.PeerGradBucket):                   # allreduce + :136-140 in one launch (K8)
            grad_norm = self._optimizer.step_peer(bucket, self._max_grad_norm)[0]
            sc = bucket.extra_reduced
        else:
            bucket.allreduce(self._group)
            if isinstance(self._optimizer, optim.FusedClipAdam):           # :136-140 in one launch (K6)
                grad_norm = self._optimizer.step_flat(bucket.grads, self._max_grad_norm)[0]
            else:
                grad_norm = bucket.clip_(self._max_grad_norm)                 # :136-137
                self._optimizer.step()                                        # :140
            sc = bucket.extra
        return {
            "train/loss": sc[ops.S_LOSS], "train/entropy": sc[ops.S_ENTROPY], "train/td": sc[ops.S_TD],
            "train/pg": sc[ops.S_PG], "train/target": sc[ops.S_TARGET],
            "train/kl": sc[ops.S_KL].clamp_min(0.), "train/ratio": sc[ops.S_RATIO],
            "train_step/grad_norm": grad_norm,
        }
