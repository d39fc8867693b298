"""Drop-ins for the rlego calls on the learner path.

* ``batched_vtrace(v_tm1, v_t, r_t, discount_t, rho_tm1)`` -- reference
  ``agents/impala/learning.py:15-26`` (``functorch.vmap(rlego.vtrace_td_error_and_advantage)``):
  ``[B, T-1]`` tensors, batch on dim 0, returns ``(adv, err, q)`` in the call-site order
  (``:150``).  Differentiable w.r.t. ``v_tm1``, ``v_t`` and ``rho_tm1`` with rlax's
  ``stop_target_gradients=True`` semantics.
* ``lambda_returns`` / ``truncated_generalized_advantage_estimation`` -- reference
  ``agents/ppo/learning.py:68`` / ``losses.py:11``, batched over columns.
"""
import torch

from . import ops


class _VTrace(torch.autograd.Function):
    @staticmethod
    def forward(ctx, v_tm1, v_t, r_t, discount_t, rho_tm1, lambda_, clip_rho, clip_pg_rho):
        adv, err, q, saved, layout = ops.vtrace_fwd(v_tm1, v_t, r_t, discount_t, rho_tm1,
                                                    lambda_, clip_rho, clip_pg_rho)
        ctx.save_for_backward(q, *saved)
        ctx.layout, ctx.lambda_, ctx.clip_pg_rho = layout, lambda_, clip_pg_rho
        return adv, err, q

    @staticmethod
    def backward(ctx, g_adv, g_err, g_q):
        q, *saved = ctx.saved_tensors
        d_vtm1, d_vt, d_rho = ops.vtrace_bwd(saved, ctx.layout, q, g_adv, g_err, g_q,
                                             ctx.lambda_, ctx.clip_pg_rho)
        return d_vtm1, d_vt, None, None, d_rho, None, None, None


def batched_vtrace(v_tm1, v_t, r_t, discount_t, rho_tm1, lambda_=1.0, clip_rho_threshold=1.0,
                   clip_pg_rho_threshold=1.0, output_order="call_site"):
    """``(pg_advantage, errors, q_estimate)``; ``output_order="rlax"`` gives
    ``(errors, pg_advantage, q_estimate)``.  Raises ``RuntimeError`` off-GPU (no fallback)."""
    adv, err, q = _VTrace.apply(v_tm1, v_t, r_t, discount_t, rho_tm1, float(lambda_),
                                float(clip_rho_threshold), float(clip_pg_rho_threshold))
    if output_order == "rlax":
        return err, adv, q
    return adv, err, q


def vtrace_td_error_and_advantage(v_tm1, v_t, r_t, discount_t, rho_tm1, lambda_=1.0, **kw):
    """Per-sequence ``[T-1]`` form (the function the reference vmaps), via the batched kernel."""
    outs = batched_vtrace(*[x.unsqueeze(0) for x in (v_tm1, v_t, r_t, discount_t, rho_tm1)],
                          lambda_=lambda_, **kw)
    return tuple(o.squeeze(0) for o in outs)


def lambda_returns(r_t, discount_t, v_t, lambda_=1.0, time_major=False):
    """Batched ``rlego.lambda_returns``: ``[B,T]`` inputs (``[T,B]`` if ``time_major``; a 1-D
    ``[T]`` rollout is accepted as in the reference).  Not differentiable (targets)."""
    if r_t.dim() == 1:
        return ops.lambda_returns(r_t[None], discount_t[None], v_t[None], float(lambda_))[0]
    if time_major:
        return ops.lambda_returns(r_t.t(), discount_t.t(), v_t.t(), float(lambda_)).t()
    return ops.lambda_returns(r_t, discount_t, v_t, float(lambda_))


def truncated_generalized_advantage_estimation(r_t, discount_t, lambda_, values, time_major=False):
    """Batched GAE; ``values`` has one more step than ``r_t`` (argument order as in rlax)."""
    if r_t.dim() == 1:
        return ops.gae(r_t[None], discount_t[None], values[None], float(lambda_))[0]
    if time_major:
        return ops.gae(r_t.t(), discount_t.t(), values.t(), float(lambda_)).t()
    return ops.gae(r_t, discount_t, values, float(lambda_))


def n_step_bootstrapped_returns(r_t, discount_t, v_t, n, lambda_t=1.0, time_major=False):
    """Batched ``rlego.n_step_bootstrapped_returns`` (reference ``agents/dqn/learning.py:62,79``); a 1-D
    ``[T]`` rollout is accepted as in the reference.  Not differentiable (targets)."""
    if r_t.dim() == 1:
        return ops.n_step_returns(r_t[None], discount_t[None], v_t[None], n, float(lambda_t))[0]
    if time_major:
        return ops.n_step_returns(r_t.t(), discount_t.t(), v_t.t(), n, float(lambda_t)).t()
    return ops.n_step_returns(r_t, discount_t, v_t, n, float(lambda_t))
