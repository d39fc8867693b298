"""``capture_graph()`` for the drop-in learners: the whole training step as ONE CUDA graph launch.

The reference's learners run one Python pass per step (``agents/impala/learning.py:119-177``,
``agents/ppo/learning.py:100-143``: RPC sample, ~180 kernel launches, host-side clip and Adam).  Here the step is
``ReplayBuffer.sample`` (K4 + K5, one launch) -> model forward -> fused loss (K1 / K3) -> model backward -> K6 (one rank)
or K8 (allreduce fused into clip + Adam over peer memory); everything that changes from step to step lives in device
memory (the sampler's draw count, Adam's step, K8's flag generation), so one capture can be replayed for ever."""
import time

import torch

from .. import ops, optim, parallel


class GraphedStepMixin:
    """Needs ``self._model, _replay_buffer, _optimizer, _batch_size, _group, _bucket, _model_push_period`` and
    ``_train_step(batch)``; counts steps in ``self._step_count``."""

    _graph = None
    _step_count = 0

    def capture_graph(self, warmup: int = 3) -> None:
        """Capture the step once; ``train_step`` then replays ONE graph launch.  Needs ``FusedClipAdam`` and the uniform
        device sampler; the replay buffer must not change SIZE between replays (a full ring may keep being appended to).
        The ``warmup`` steps before the capture are real training steps.  Every rank must call it at the same point."""
        if self._graph is not None:
            return
        if not isinstance(self._optimizer, optim.FusedClipAdam):
            raise RuntimeError("capture_graph needs impala_b200.optim.FusedClipAdam (a launch-argument step number cannot be replayed)")
        from .. import replay as _replay
        sampler = self._replay_buffer.sampler
        if type(sampler) is not _replay.UniformSampler:
            raise RuntimeError("capture_graph supports the uniform device sampler")
        dev = self.device()
        if self._bucket is None:
            self._bucket = parallel.make_bucket(self._model.parameters(), self._group, True)
        if not isinstance(self._bucket, parallel.PeerGradBucket) and parallel.world(self._group)[1] > 1:
            raise RuntimeError("capture_graph across ranks needs the peer-memory gradient bucket (K8); an NCCL allreduce is not captured here")
        self._optimizer.capturable_(self._bucket if isinstance(self._bucket, parallel.PeerGradBucket) else None)
        sampler.capturable_(dev)
        self.profile = False
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            ops.workspace(dev)                      # the capture stream's loss workspace must exist before capture
            for _ in range(max(1, warmup)):         # cuDNN plans, allocator pools, lazy module loads
                self._step_body()                   # a real training step: it counts towards the model-push period
                self._push_model()
        side.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            self._graph_metrics = self._step_body()
        torch.cuda.current_stream(dev).wait_stream(side)
        self._graph = graph

    def _step_body(self):
        _, batch, _ = self._replay_buffer.sample(self._batch_size)
        self._batch_shape = tuple(batch[0].shape)
        return self._train_step(batch)

    def _push_model(self) -> float:
        """``model.push()`` every ``model_push_period`` steps (reference ``agents/impala/learning.py:127-131``,
        ``agents/ppo/learning.py:108-112``); returns the seconds it took."""
        self._step_count += 1
        if self._step_count % self._model_push_period == 0 and hasattr(self._model, "push"):
            start = time.perf_counter()
            self._model.push()
            return time.perf_counter() - start
        return 0

    def _replay_step(self):
        """One graph launch; ``(metrics, seconds of the launch call)``.  The metric tensors are static device tensors that
        every replay rewrites."""
        t0 = time.perf_counter()
        self._graph.replay()
        return dict(self._graph_metrics), time.perf_counter() - t0
