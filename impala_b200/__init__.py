"""impala_b200 -- B200-native learner hot path for d3sm0/impala (drop-in).

Host side (PyTorch: device memory, streams, autograd glue, torch.distributed) over a
C-ABI CUDA library for sm_100a (``include/b200rl.h``, ``impala_b200/csrc``).
Mirrors the reference's operator API for this path only:

* ``impala_b200.vtrace.batched_vtrace``          <- ``agents/impala/learning.py:15-26``
* ``impala_b200.losses.{ppo_loss, impala_loss}`` <- ``losses.py:92,131``
* ``impala_b200.agents.impala.ImpalaLearner``    <- ``agents/impala/learning.py:86-177``
* ``impala_b200.agents.ppo.PPOLearner``          <- ``agents/ppo/learning.py:75-143``
* ``impala_b200.replay.ReplayBuffer`` et al.     <- rlmeta ``ReplayBuffer/CircularBuffer/UniformSampler``

There is no CPU fallback: every op raises ``RuntimeError`` without the CUDA library
or a CUDA device.
"""
from . import _cabi  # noqa: F401

__version__ = "0.1.0"
