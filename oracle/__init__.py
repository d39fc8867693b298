"""CPU oracle for the learner hot path of d3sm0/impala -- TEST INFRASTRUCTURE ONLY.

This package is the *checker*, never the product: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it.  Nothing under ``impala_b200/`` imports it.

PARITY UNPINNED for the rlego / rlmeta arithmetic (pinned for the rest, see the end of this note).  The reference (``/root/reference``) holds no tests, no golden
vectors and no fixtures, and the arithmetic of its hot path lives
in third-party packages that are neither vendored nor installed nor pinned:

* ``rlego``  -- ``git+https://github.com/d3sm0/RLego.git`` HEAD, no commit hash
  (reference ``requirements.txt:7``).  Restated here from the published
  algorithms its function names mirror one-for-one (DeepMind rlax
  ``vtrace_td_error_and_advantage`` / ``lambda_returns`` /
  ``truncated_generalized_advantage_estimation``; Espeholt et al. 2018 sec 4.1;
  Schulman et al. 2016), anchored on the reference's own call sites
  (``agents/impala/learning.py:18,150-153``, ``losses.py:106``,
  ``agents/ppo/learning.py:68``).
* ``rlmeta`` -- facebookresearch/rlmeta HEAD, unpinned (reference
  ``Dockerfile:43-47``): ``ReplayBuffer`` / ``CircularBuffer`` /
  ``UniformSampler``.  Its RNG stream cannot be reproduced (C++ extension
  absent), so the sampler is *specified* here (Philox4x32-10, pinned by the
  Random123 known-answer vectors) and the CUDA path must match it bit for bit.

What *is* pinned: the torch arithmetic the reference itself calls
(``torch.distributions.Categorical`` / ``kl_divergence`` are imported from the
installed torch, not restated), and -- by fixtures produced by RUNNING the
reference's own files with import shims for the absent third-party names
(``tests/golden/make_golden_from_reference.py`` -> ``tests/golden/ref_*.npz``;
checked in ``tests/test_reference_golden.py``) -- ``losses.ppo_loss``,
``quantile_regression_loss``, the whole of ``ImpalaLearner._train_step`` around
the rlego call (slices, means, signs, no-detach, clip, Adam) and
``losses.impala_loss`` as committed.  The rlego recursions themselves cannot be
checked against rlego (unobtainable): they are covered by hand-derived known-answer
cases and -- an EXTERNAL pin of the algorithms they restate -- by the known answers
of rlax's own unit tests for V-trace (TD errors, policy-gradient advantages),
lambda-returns and GAE (``tests/rlax_vectors.py``, which states how those vectors
got here; ``tests/test_oracle.py::test_rlax_published_vectors``).  That rlego
computes what rlax computes remains the assumption its function names invite.
"""
