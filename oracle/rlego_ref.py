"""Restatement of the ``rlego`` functions the reference calls (TEST ORACLE, CPU).

``rlego`` is absent and unpinned (see ``oracle/__init__.py``): PARITY UNPINNED against rlego itself; the rlax algorithms
restated here reproduce the known answers of rlax's own unit tests (``tests/rlax_vectors.py``).
Per-sequence ``[T]`` functions with a Python loop over time, exactly the shape
the reference feeds to ``functorch.vmap`` (``agents/impala/learning.py:18``).
Algorithms follow DeepMind rlax (the package rlego mirrors name-for-name).

Every unverifiable semantic is a named switch:
``output_order``, ``stop_target_gradients``, clip thresholds, ``lambda_``.
"""
import torch


def vtrace(v_tm1, v_t, r_t, discount_t, rho_tm1, lambda_=1.0,
           clip_rho_threshold=1.0, stop_target_gradients=True):
    """V-trace errors ``vs_t - v_t`` (Espeholt et al. 2018, eq. 1).

    Call sites: reference ``agents/impala/learning.py:150-153`` (through
    ``vtrace_td_error_and_advantage``).
    """
    c_tm1 = torch.clamp(rho_tm1, max=1.0) * lambda_
    clipped_rhos_tm1 = torch.clamp(rho_tm1, max=clip_rho_threshold)
    td_errors = clipped_rhos_tm1 * (r_t + discount_t * v_t - v_tm1)
    err = torch.zeros_like(td_errors[0])
    errors = []
    for i in reversed(range(v_t.shape[0])):
        err = td_errors[i] + discount_t[i] * c_tm1[i] * err
        errors.insert(0, err)
    errors = torch.stack(errors)
    targets_tm1 = errors + v_tm1
    if stop_target_gradients:
        targets_tm1 = targets_tm1.detach()
    return targets_tm1 - v_tm1


def vtrace_td_error_and_advantage(v_tm1, v_t, r_t, discount_t, rho_tm1,
                                  lambda_=1.0, clip_rho_threshold=1.0,
                                  clip_pg_rho_threshold=1.0,
                                  stop_target_gradients=True,
                                  output_order="call_site"):
    """Returns ``(pg_advantage, errors, q_estimate)`` -- the order the reference
    binds at ``agents/impala/learning.py:150`` (``adv, err, _``) and
    ``losses.py:38,73,106``.  ``output_order="rlax"`` gives rlax's NamedTuple
    order ``(errors, pg_advantage, q_estimate)``.
    """
    errors = vtrace(v_tm1, v_t, r_t, discount_t, rho_tm1, lambda_,
                    clip_rho_threshold, stop_target_gradients)
    targets_tm1 = errors + v_tm1
    q_bootstrap = torch.cat([
        lambda_ * targets_tm1[1:] + (1 - lambda_) * v_tm1[1:],
        v_t[-1:],
    ], dim=0)
    q_estimate = r_t + discount_t * q_bootstrap
    clipped_pg_rho_tm1 = torch.clamp(rho_tm1, max=clip_pg_rho_threshold)
    pg_advantages = clipped_pg_rho_tm1 * (q_estimate - v_tm1)
    if output_order == "rlax":
        return errors, pg_advantages, q_estimate
    return pg_advantages, errors, q_estimate


def lambda_returns(r_t, discount_t, v_t, lambda_=1.0, stop_target_gradients=False):
    """``G_t = r_t + g_t ((1-l) v_{t+1} + l G_{t+1})``, bootstrap ``G_T = v_T``.

    Call site: reference ``agents/ppo/learning.py:68`` with
    ``(r[:-1], discount[:-1], v[1:], 0.95)``.
    """
    g = v_t[-1]
    returns = []
    for i in reversed(range(v_t.shape[0])):
        g = r_t[i] + discount_t[i] * ((1 - lambda_) * v_t[i] + lambda_ * g)
        returns.insert(0, g)
    returns = torch.stack(returns)
    return returns.detach() if stop_target_gradients else returns


def truncated_generalized_advantage_estimation(r_t, discount_t, lambda_, values,
                                               stop_target_gradients=False):
    """GAE (Schulman et al. 2016): ``values`` is ``[T+1]``, result ``[T]``.

    Named only in a comment in the reference (``losses.py:11``).
    """
    delta_t = r_t + discount_t * values[1:] - values[:-1]
    gae = torch.zeros_like(delta_t[0])
    adv = []
    for t in reversed(range(delta_t.shape[0])):
        gae = delta_t[t] + lambda_ * discount_t[t] * gae
        adv.insert(0, gae)
    adv = torch.stack(adv)
    return adv.detach() if stop_target_gradients else adv


def n_step_bootstrapped_returns(r_t, discount_t, v_t, n, lambda_t=1.0,
                                stop_target_gradients=False):
    """rlax ``n_step_bootstrapped_returns`` (reference ``agents/dqn/learning.py:62,79``;
    SURVEY sec 8f row N4 -- "next", restated for completeness)."""
    seq_len = r_t.shape[0]
    pad = torch.ones(n - 1, dtype=v_t.dtype) * v_t[-1]
    targets = torch.cat([v_t[n - 1:], pad])
    r_pad = torch.cat([r_t, torch.zeros(n - 1, dtype=r_t.dtype)])
    d_pad = torch.cat([discount_t, torch.ones(n - 1, dtype=discount_t.dtype)])
    v_pad = torch.cat([v_t, pad])
    for i in reversed(range(n)):
        r_ = r_pad[i:i + seq_len]
        d_ = d_pad[i:i + seq_len]
        v_ = v_pad[i:i + seq_len]
        targets = r_ + d_ * ((1.0 - lambda_t) * v_ + lambda_t * targets)
    return targets.detach() if stop_target_gradients else targets


def batched(fn, **kw):
    """``functorch.vmap(fn)`` over dim 0, as ``agents/impala/learning.py:18``."""
    return torch.func.vmap(lambda *a: fn(*a, **kw))


def batched_loop(fn, *batch, **kw):
    """The reference's fallback: Python loop over B then stack
    (``agents/impala/learning.py:20-26``)."""
    outs = [fn(*[x[b] for x in batch], **kw) for b in range(batch[0].shape[0])]
    if isinstance(outs[0], tuple):
        return tuple(torch.stack(o) for o in zip(*outs))
    return torch.stack(outs)
