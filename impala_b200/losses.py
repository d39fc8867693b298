"""Drop-in for the reference's ``losses.py`` on the learner path (fused CUDA forward+backward).

``ppo_loss(model, batch, entropy_cost, clip_coeff)`` (reference ``losses.py:131``) and
``impala_loss(batch, model, lambda_, entropy_cost)`` (``losses.py:92``; note the argument order)
keep their signatures, return ``(loss, metrics)`` with the same metric keys, and ``loss.backward()``
reaches the model parameters.  The loss value, every metric and d loss/d(model outputs) come out
of ONE kernel launch; autograd only carries the precomputed gradients into the model.
"""
import torch

from . import ops
from . import vtrace as _vtrace

METRIC_SLOTS = (("train/loss", ops.S_LOSS), ("train/entropy", ops.S_ENTROPY), ("train/td", ops.S_TD),
                ("train/pg", ops.S_PG), ("train/kl", ops.S_KL), ("train/ratio", ops.S_RATIO))


_warned = set()


def _as_input(x, name, dtype=torch.float32):
    """The drop-in functions accept whatever the reference's torch code accepts (a sliced head output, a float64
    reward, an int32 action ...); the kernels take contiguous float32 / int64 and ``ops`` raises on anything else.  The
    conversion happens HERE, explicitly, and is announced once per (argument, reason): it is a copy the caller can
    remove by producing the right layout."""
    if x.dtype == dtype and x.is_contiguous():
        return x
    key = (name, str(x.dtype), x.is_contiguous())
    if key not in _warned:
        _warned.add(key)
        import warnings
        warnings.warn("impala_b200.losses: %s arrives as %s%s and is copied to contiguous %s before the fused kernel"
                      % (name, x.dtype, "" if x.is_contiguous() else " (non-contiguous)", dtype))
    return x.to(dtype).contiguous()


class _FusedLoss(torch.autograd.Function):
    """loss = f(outputs...) whose gradients were already produced by the fused kernel."""

    @staticmethod
    def forward(ctx, loss, n_out, *args):
        outs, grads = args[:n_out], args[n_out:]
        ctx.save_for_backward(*grads)
        ctx.shapes = [o.shape for o in outs]
        return loss.clone()

    @staticmethod
    def backward(ctx, g):
        grads = ctx.saved_tensors
        return (None, None) + tuple((d * g).reshape(s) for d, s in zip(grads, ctx.shapes)) + (None,) * len(grads)


def _attach(loss, outputs, grads):
    if not any(o.requires_grad for o in outputs):
        return loss
    return _FusedLoss.apply(loss, len(outputs), *outputs, *grads)


def impala_loss_from_outputs(logits, values, a, r, discount_t, pi_ref, entropy_coeff=0.01, lambda_=1.0,
                             clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0, detach_adv=False,
                             detach_rho=False, time_major=False, n_global_b=None):
    """The loss block of ``ImpalaLearner._train_step`` (reference ``agents/impala/learning.py:146-170``)
    on model outputs ``logits [B,T,A]``, ``values [B,T]``.  Returns ``(loss, metrics)``."""
    flags = (ops.DETACH_ADV if detach_adv else 0) | (ops.DETACH_RHO if detach_rho else 0)
    scalars, dlogits, dvalues = ops.impala_loss_fwdbwd(
        _as_input(logits, "logits"), _as_input(values, "values"), _as_input(a, "a", torch.int64), _as_input(r, "r"),
        _as_input(discount_t, "discount_t"), _as_input(pi_ref, "pi_ref"), entropy_coeff, lambda_, clip_rho_threshold,
        clip_pg_rho_threshold, flags, n_global_b, time_major)
    loss = _attach(scalars[ops.S_LOSS], (logits, values), (dlogits, dvalues))
    metrics = {k: scalars[i] for k, i in METRIC_SLOTS}
    return loss, metrics


def ppo_loss_from_outputs(logits, v, a, v_target, pi_ref, entropy_cost=0.01, clip_coeff=0.1, n_global=None):
    """Reference ``losses.py:134-155`` on model outputs ``logits [N,A]``, ``v [N,1]``."""
    scalars, dlogits, dv = ops.ppo_loss_fwdbwd(_as_input(logits, "logits"), _as_input(v, "v"), _as_input(a, "a", torch.int64),
                                               _as_input(v_target, "v_target"), _as_input(pi_ref, "pi_ref"),
                                               entropy_cost, clip_coeff, n_global)
    loss = _attach(scalars[ops.S_LOSS], (logits, v), (dlogits, dv))
    metrics = {k: scalars[i] for k, i in METRIC_SLOTS}
    metrics["train/target"] = scalars[ops.S_TARGET]
    metrics["train/kl"] = scalars[ops.S_KL].clamp_min(0.)  # losses.py:144
    order = ("train/loss", "train/entropy", "train/td", "train/pg", "train/target", "train/kl", "train/ratio")
    return loss, {k: metrics[k] for k in order}


def ppo_loss(model, batch, entropy_cost=0.01, clip_coeff=0.1):
    """Drop-in for reference ``losses.ppo_loss`` (``losses.py:131-155``)."""
    s, a, v_target, pi_ref = batch
    pi_tm1, v_tm1 = model(s)
    return ppo_loss_from_outputs(pi_tm1, v_tm1, a, v_target, pi_ref, entropy_cost, clip_coeff)


def ppo_is_loss_from_outputs(logits, v_tm1, v_boot, a, r, d, pi_ref, entropy_cost=0.01, clip_coeff=0.1,
                             lambda_=0.95, gamma=0.99, n_global_b=None):
    """Reference ``losses.py:64-89`` on model outputs ``logits [B,T,A]``, ``v_tm1 [B,T]`` and the bootstrap
    value ``v_boot [B]`` of ``s1[:, -1]``."""
    scalars, dlogits, dv = ops.ppo_is_loss_fwdbwd(_as_input(logits, "logits"), _as_input(v_tm1, "v_tm1"), _as_input(v_boot, "v_boot"),
                                                  _as_input(a, "a", torch.int64), _as_input(r, "r"), d, _as_input(pi_ref, "pi_ref"),
                                                  entropy_cost, clip_coeff, lambda_, gamma, n_global_b=n_global_b)
    loss = _attach(scalars[ops.S_LOSS], (logits, v_tm1), (dlogits, dv))
    metrics = {k: scalars[i] for k, i in METRIC_SLOTS}
    metrics["train/kl"] = scalars[ops.S_KL].clamp_min(0.)  # losses.py:79
    return loss, metrics


def ppo_loss_is(batch, model, entropy_cost=0.01, clip_coeff=0.1, lambda_=0.95, gamma=0.99):
    """Drop-in for reference ``losses.ppo_loss_is`` (``losses.py:62-89``): same signature, ``(loss, metrics)``
    return and metric keys.  The committed function cannot run (``functorch`` is not imported, ``batched_vtrace``
    is undefined there); this computes what it spells out.  ``functorch.vmap(model)(s)`` becomes one model call
    on the flattened ``[B*T,...]`` batch; the bootstrap call on ``s1[:, -1:]`` runs under ``no_grad`` (``:70-72``)."""
    s, a, r, s1, d, pi_ref = batch
    B, T = s.shape[0], s.shape[1]
    pi, v = model(s.flatten(0, 1))
    pi = pi.reshape(B, T, -1)
    v = v.reshape(B, T)
    with torch.no_grad():
        _, v_boot = model(s1[:, -1])
    sq = lambda x: x.squeeze(-1) if x.dim() == 3 else x  # noqa: E731
    return ppo_is_loss_from_outputs(pi, v, v_boot.reshape(B), sq(a), sq(r), sq(d), pi_ref, entropy_cost, clip_coeff,
                                    lambda_, gamma)


def impala_loss(batch, model, lambda_=1., entropy_cost=0.01, as_written=False):
    """Drop-in for reference ``losses.impala_loss`` (``losses.py:92-128``).

    The committed reference function is work in progress and has no caller (raw logits instead of
    log-probs, ratio = 1, same-step bootstrap, entropy disabled).  By default this computes the loss
    the IMPALA learner actually trains with (``agents/impala/learning.py:146-159``) on a ``[B,T,...]``
    batch; ``as_written=True`` computes the committed function literally (``_impala_loss_as_written``)."""
    if as_written:
        return _impala_loss_as_written(batch, model, lambda_)
    s, a, r, discount_t, pi_ref = batch
    B, T = s.shape[0], s.shape[1]
    pi, values = model(s.flatten(0, 1))
    pi = pi.reshape(B, T, -1)
    values = values.reshape(B, T)
    sq = lambda x: x.squeeze(-1) if x.dim() == 3 else x  # noqa: E731
    return impala_loss_from_outputs(pi, values, sq(a), sq(r), sq(discount_t), pi_ref, entropy_cost, lambda_)


def _impala_loss_as_written(batch, model, lambda_):
    """Reference ``losses.impala_loss`` literally (``losses.py:92-128``): ONE un-batched sequence ``[N,...]``, raw
    logits gathered instead of log-probs (``:116``), ``ratio = ones`` (``:99``), the bootstrap is the same-step
    detached value (``:105``), ``td = 0.5 mean err^2`` (``:109``), ``pg = -sum`` (``:116``), entropy disabled
    (``:120``).  The V-trace call (``:106-108``) is the CUDA kernel behind ``vtrace.batched_vtrace`` (differentiable:
    ``adv`` is not detached here either); the gather and the two reductions around it are the three torch ops the
    reference itself spells out."""
    s, a, r, discount_t, pi_ref = batch
    pi_tm1, v_tm1 = model(s)
    a = a.squeeze(dim=-1)
    ratio = torch.ones_like(r).squeeze().float()
    v = v_tm1.squeeze(-1)
    adv, err, _ = _vtrace.vtrace_td_error_and_advantage(v, v.detach(), r.squeeze(dim=-1), discount_t.squeeze(dim=-1),
                                                        ratio, lambda_=lambda_)
    td_loss = 0.5 * err.pow(2).mean()
    pg_loss = -(pi_tm1.gather(1, a.unsqueeze(dim=-1)).squeeze(dim=-1) * adv).sum()
    loss = pg_loss + td_loss
    return loss, {"train/loss": loss.detach(), "train/td": td_loss.detach(), "train/pg": pg_loss.detach(),
                  "train/ratio": ratio.mean().detach()}


class _QuantileHuber(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dist_src, tau_src, dist_target, huber_param):
        loss, d_src = ops.quantile_huber_fwdbwd(dist_src, tau_src, dist_target, huber_param)
        ctx.save_for_backward(d_src)
        ctx.shape = dist_src.shape
        return loss

    @staticmethod
    def backward(ctx, g):
        (d_src,) = ctx.saved_tensors
        return (d_src * g.reshape(-1, 1)).reshape(ctx.shape), None, None, None


def quantile_regression_loss(dist_src, tau_src, dist_target, huber_param: float = 1.):
    """Drop-in for reference ``models/quantile_layers.py:83-93`` (per-sample loss ``[B]``, differentiable w.r.t.
    ``dist_src``; ``tau_src`` enters only through the detached weights, ``dist_target`` is a target)."""
    return _QuantileHuber.apply(dist_src, tau_src, dist_target, float(huber_param))


def per_td_loss(q, action, target, probabilities, importance_sampling_exponent: float):
    """The loss block of ``ApexLearner._train_step`` (reference ``agents/dqn/learning.py:161-172``) on the Q network's
    output ``q [N,A]``: ``loss = 0.5 * mean((target - q.gather(1, action))^2 * w)``, ``w`` = the prioritized sampler's
    probabilities turned into max-normalised importance weights.  Returns ``(loss, priorities, metrics)``:
    ``loss.backward()`` reaches the Q network, ``priorities = |target - q|`` go back to ``ReplayBuffer.update`` (``:184``),
    ``metrics`` carry the reference's ``train_step/*`` keys as device scalars.  One kernel launch (N3b)."""
    scalars, dq, priorities, _ = ops.per_td_loss_fwdbwd(_as_input(q, "q"), _as_input(action, "action", torch.int64),
                                                        _as_input(target, "target"), _as_input(probabilities, "probabilities"),
                                                        importance_sampling_exponent)
    loss = _attach(scalars[0], (q,), (dq,))
    metrics = {"train_step/q": scalars[1], "train_step/target": scalars[2], "train_step/priorities": scalars[3],
               "train_step/loss": scalars[0]}
    return loss, priorities, metrics
