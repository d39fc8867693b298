"""Restatement of the reference's loss code (TEST ORACLE, CPU torch, differentiable).

PARITY UNPINNED at the rlego boundary (see ``oracle/__init__.py``).  The
``torch.distributions`` arithmetic is the installed torch's own, as in the
reference.  Each function takes the *model outputs* (logits, values) instead of
``model`` + observations so that the CUDA path can be compared on identical
tensors; ``*_with_model`` wrappers keep the reference signatures.
"""
import torch
from torch.distributions import Categorical, kl_divergence

from . import rlego_ref


def impala_train_loss(logits, values, a, r, discount_t, pi_ref_logits,
                      entropy_coeff=0.01, lambda_=1.0, clip_rho_threshold=1.0,
                      clip_pg_rho_threshold=1.0, detach_adv=False, detach_rho=False,
                      use_vmap=True):
    """Reference ``ImpalaLearner._train_step`` lines ``agents/impala/learning.py:146-170``.

    Batch-major: ``logits, pi_ref_logits [B,T,A]``, ``values, r, discount_t [B,T]``,
    ``a [B,T] int64``.  Returns ``(loss, metrics)`` with the reference's metric keys
    (minus ``train/grad_norm`` which belongs to the optimiser step).
    ``detach_adv`` / ``detach_rho`` default False = call-site truth (``:148,155``).
    """
    pi = Categorical(logits=logits)                                   # :146
    pi_ref = Categorical(logits=pi_ref_logits)                        # :147
    rho_tm1 = torch.exp(pi.log_prob(a) - pi_ref.log_prob(a))          # :148
    rho_in = rho_tm1.detach() if detach_rho else rho_tm1
    kw = dict(lambda_=lambda_, clip_rho_threshold=clip_rho_threshold,
              clip_pg_rho_threshold=clip_pg_rho_threshold)
    args = (values[:, :-1], values[:, 1:], r[:, :-1], discount_t[:, :-1], rho_in[:, :-1])
    if use_vmap:                                                      # :18
        adv, err, _ = rlego_ref.batched(rlego_ref.vtrace_td_error_and_advantage, **kw)(*args)
    else:                                                             # :20-26
        adv, err, _ = rlego_ref.batched_loop(rlego_ref.vtrace_td_error_and_advantage, *args, **kw)
    if detach_adv:
        adv = adv.detach()
    pg_loss = (pi.log_prob(a)[:, :-1] * adv).mean()                   # :155
    value_loss = err.pow(2).mean()                                    # :156
    entropy_loss = pi.entropy().mean()                                # :157
    loss = -pg_loss + value_loss - entropy_coeff * entropy_loss       # :159
    metrics = {
        "train/loss": loss.detach(),
        "train/entropy": entropy_loss.detach(),
        "train/td": value_loss.detach(),
        "train/pg": pg_loss.detach(),
        "train/kl": kl_divergence(pi, pi_ref).mean().detach(),        # :167
        "train/ratio": rho_tm1.mean().detach(),                       # :168
    }
    return loss, metrics


def ppo_loss_from_outputs(logits, v, a, v_target, pi_ref_logits,
                          entropy_cost=0.01, clip_coeff=0.1):
    """Reference ``losses.ppo_loss`` lines ``losses.py:134-155`` after ``model(s)``.

    ``logits, pi_ref_logits [N,A]``, ``v [N,1]`` or ``[N]``, ``a [N] int64``,
    ``v_target [N]``.
    """
    pi_tm1 = Categorical(logits=logits)
    pi_ref = Categorical(logits=pi_ref_logits)
    ratio = torch.exp(pi_tm1.log_prob(a) - pi_ref.log_prob(a))        # :136
    adv = v_target - (v.squeeze(-1) if v.dim() > v_target.dim() else v)  # :137
    td_loss = 0.5 * adv.pow(2).mean()                                 # :138
    adv = adv.detach()                                                # :139
    pg_loss_1 = -(adv * ratio)                                        # :140
    pg_loss_2 = -torch.clamp(ratio, 1 - clip_coeff, 1 + clip_coeff) * adv  # :141
    pg_loss = torch.max(pg_loss_1, pg_loss_2).mean()                  # :142
    kl = kl_divergence(pi_tm1, pi_ref).mean().clamp_min(0.)           # :144
    entropy = pi_tm1.entropy().mean()                                 # :145
    loss = pg_loss + td_loss - entropy_cost * entropy                 # :146
    return loss, {
        "train/loss": loss.detach(),
        "train/entropy": entropy.detach(),
        "train/td": td_loss.detach(),
        "train/pg": pg_loss.detach(),
        "train/target": v_target.mean().detach(),
        "train/kl": kl.detach(),
        "train/ratio": ratio.mean().detach(),
    }


def ppo_loss(model, batch, entropy_cost=0.01, clip_coeff=0.1):
    """Reference signature, ``losses.py:131``."""
    s, a, v_target, pi_ref = batch
    logits, v = model(s)
    return ppo_loss_from_outputs(logits, v, a, v_target, pi_ref, entropy_cost, clip_coeff)


def ppo_is_loss_from_outputs(logits, v_tm1, v_boot, a, r, d, pi_ref_logits, entropy_cost=0.01,
                             clip_coeff=0.1, lambda_=0.95, gamma=0.99):
    """Reference ``losses.ppo_loss_is`` lines ``losses.py:64-89`` after the two model calls.  The
    committed function is not executable (``functorch`` not imported ``:1,66``, ``batched_vtrace``
    undefined ``:10,73``, ``1 - d`` on a bool tensor); this restates what it spells out, with
    ``batched_vtrace`` = the per-rollout loop over ``rlego_ref.vtrace_td_error_and_advantage``.

    ``logits, pi_ref_logits [B,T,A]``, ``v_tm1 [B,T]``, ``v_boot [B]`` (value of ``s1[:, -1]``, no
    gradient), ``a [B,T] int64``, ``r [B,T]``, ``d [B,T]`` done flags.
    """
    df = d.to(v_tm1.dtype)
    mask = (1 - df).roll(shifts=1, dims=1)                            # :64
    discount_t = (1 - df) * gamma                                     # :65
    pi_tm1 = Categorical(logits=logits)
    pi_ref = Categorical(logits=pi_ref_logits)
    ratio = torch.exp(pi_tm1.log_prob(a) - pi_ref.log_prob(a))        # :69
    with torch.no_grad():                                             # :70-72
        v_t = torch.cat([v_tm1[:, 1:], v_boot.reshape(-1, 1)], dim=1)
    adv, err, _ = rlego_ref.batched_loop(rlego_ref.vtrace_td_error_and_advantage, v_tm1, v_t, r,
                                         discount_t, ratio.detach(), lambda_=lambda_)   # :73
    td_loss = 0.5 * (mask * err).pow(2).mean()                        # :74
    adv = (adv * mask).detach()                                       # :75
    pg_loss_1 = -(adv * ratio)                                        # :76
    pg_loss_2 = -torch.clamp(ratio, 1 - clip_coeff, 1 + clip_coeff) * adv  # :77
    pg_loss = torch.max(pg_loss_1, pg_loss_2).mean()                  # :78
    kl = kl_divergence(pi_tm1, pi_ref).mean().clamp_min(0.)           # :79
    entropy = pi_tm1.entropy().mean()                                 # :80
    loss = pg_loss + td_loss - entropy_cost * entropy                 # :81
    return loss, {
        "train/loss": loss.detach(),
        "train/entropy": entropy.detach(),
        "train/td": td_loss.detach(),
        "train/pg": pg_loss.detach(),
        "train/kl": kl.detach(),
        "train/ratio": ratio.mean().detach(),
    }


def ppo_loss_is(batch, model, entropy_cost=0.01, clip_coeff=0.1, lambda_=0.95, gamma=0.99):
    """Reference signature, ``losses.py:62``; ``functorch.vmap(model)(s)`` = the model applied per rollout."""
    s, a, r, s1, d, pi_ref = batch
    B, T = s.shape[0], s.shape[1]
    logits, v = model(s.flatten(0, 1))
    with torch.no_grad():
        _, v_boot = model(s1[:, -1])
    sq = lambda x: x.squeeze(-1) if x.dim() == 3 else x  # noqa: E731
    return ppo_is_loss_from_outputs(logits.reshape(B, T, -1), v.reshape(B, T), v_boot.reshape(B), sq(a), sq(r),
                                    sq(d), pi_ref, entropy_cost, clip_coeff, lambda_, gamma)


def impala_loss_from_outputs(pi_tm1, v_tm1, a, r, discount_t, lambda_=1.0):
    """Reference ``losses.impala_loss`` AS WRITTEN (``losses.py:98-128``; WIP, no caller):
    raw logits gathered (not log-probs), ``ratio = ones``, bootstrap = same-time detached
    value, one un-batched sequence ``[N,...]``, ``pg = -sum``, entropy disabled.
    """
    a = a.squeeze(dim=-1)                                             # :98
    ratio = torch.ones_like(r).squeeze().float()                      # :99
    v_t = v_tm1.squeeze(-1).detach()                                  # :105
    adv, err, _ = rlego_ref.vtrace_td_error_and_advantage(            # :106-108
        v_tm1.squeeze(-1), v_t, r.squeeze(dim=-1), discount_t.squeeze(dim=-1),
        ratio.detach(), lambda_=lambda_)
    td_loss = 0.5 * err.pow(2).mean()                                 # :109
    pg_loss = -(pi_tm1.gather(1, a.unsqueeze(dim=-1)).squeeze(dim=-1) * adv).sum()  # :116
    loss = pg_loss + td_loss                                          # :120
    return loss, {
        "train/loss": loss.detach(),
        "train/td": td_loss.detach(),
        "train/pg": pg_loss.detach(),
        "train/ratio": ratio.mean().detach(),
    }


def impala_loss(batch, model, lambda_=1., entropy_cost=0.01):
    """Reference signature, ``losses.py:92`` (note: batch first, unlike ``ppo_loss``)."""
    s, a, r, discount_t, pi_ref = batch
    pi_tm1, v_tm1 = model(s)
    return impala_loss_from_outputs(pi_tm1, v_tm1, a, r, discount_t, lambda_)


def ppo_make_targets(r, done, v, gamma=0.99, lambda_=0.95):
    """Reference ``PPOActorRemote._make_replay`` lines ``agents/ppo/learning.py:67-68``
    with the evident intent ``values -> v`` (the committed line raises NameError).
    Per rollout ``[T]`` -> ``[T-1]``."""
    discount_t = torch.logical_not(done) * gamma
    return rlego_ref.lambda_returns(r[:-1], discount_t[:-1], v[1:], lambda_)


def quantile_regression_loss(dist_src, tau_src, dist_target, huber_param=1.):
    """Reference ``models/quantile_layers.py:83-93`` restated line by line."""
    delta = dist_target.unsqueeze(1) - dist_src.unsqueeze(2)
    weight = torch.abs(tau_src.unsqueeze(2) - (delta < 0.).to(torch.float32)).detach()
    if huber_param == 0:
        delta = torch.abs(delta)
    else:
        abs_delta = torch.abs(delta)
        quadratic = torch.clamp_max(abs_delta, huber_param)
        delta = 0.5 * quadratic.pow(2) + huber_param * (abs_delta - quadratic)
    return (weight * delta).sum(1).mean(-1)


def act_epilogue(logits, deterministic, seed, stream_id, draw_offset):
    """Specification of the fused actor epilogue (reference ``models/distributed_models.py:26-31``
    draws with ``torch.multinomial`` whose stream cannot be reproduced -- PARITY UNPINNED, distribution identical):
    returns ``(action [n], lo [n], hi [n])`` where ``lo <= u*total < hi`` are the cumulative-sum brackets of the
    chosen action, so a GPU result can be accepted when its own bracket contains the same target up to rounding."""
    import numpy as np
    from . import replay_ref
    x = logits.double().numpy()
    n, A = x.shape
    g = np.arange(n, dtype=np.uint64) + np.uint64(draw_offset)
    o0, _, _, _ = replay_ref.philox4x32_10_vec(g & np.uint64(0xFFFFFFFF), g >> np.uint64(32),
                                                np.full(n, stream_id, dtype=np.uint64), np.ones(n, dtype=np.uint64),
                                                seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    u = (o0 >> np.uint64(8)).astype(np.float64) * 2.0 ** -24
    e = np.exp(x - x.max(1, keepdims=True))
    cum = np.cumsum(e, 1)
    target = u * cum[:, -1]
    a = np.minimum((cum > target[:, None]).argmax(1), A - 1)
    det = np.broadcast_to(np.asarray(deterministic, dtype=bool).reshape(-1), (n,))
    a = np.where(det, x.argmax(1), a)
    return a, u, cum / cum[:, -1:]


def apex_td_loss(q_all, action, target, probabilities, importance_sampling_exponent):
    """Reference ``ApexLearner._train_step`` lines ``agents/dqn/learning.py:161-172,188-191`` on the Q network's output:
    ``(loss, priorities, metrics)``."""
    q = q_all.gather(1, action.unsqueeze(-1)).squeeze(-1)                       # :161-162
    probabilities = probabilities.to(dtype=q.dtype)                             # :164
    weight = probabilities.pow(-importance_sampling_exponent)                   # :165
    weight = weight / weight.max()                                              # :166 (div_ in place there)
    err = target - q                                                            # :167
    loss = 0.5 * (err.pow(2) * weight).mean()                                   # :168
    priorities = err.detach().abs()                                             # :172
    metrics = {"train_step/q": q.mean().detach(), "train_step/target": target.mean(),
               "train_step/priorities": priorities.mean(), "train_step/loss": loss.detach()}   # :188-191
    return loss, priorities, metrics
