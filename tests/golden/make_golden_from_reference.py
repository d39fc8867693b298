"""Golden vectors produced by RUNNING THE REFERENCE'S OWN CODE in this container (not the oracle).

    python tests/golden/make_golden_from_reference.py          # needs /root/reference (read-only)

The reference's modules cannot be imported as they stand: ``rlego``, ``rlmeta``, ``moolib``, ``envpool`` are
absent.  This script therefore installs *import shims* for exactly those third-party names in
``sys.modules`` and then executes the UNMODIFIED reference files where they lie:

=====================  =====================================================  ==========================================
fixture                reference code that ran                                 third-party arithmetic involved
=====================  =====================================================  ==========================================
``ref_ppo_loss``       ``losses.ppo_loss`` (``losses.py:131-155``)             none -> **pinned by the reference itself**
``ref_quantile``       ``models.quantile_layers.quantile_regression_loss``     none -> **pinned by the reference itself**
                       (``models/quantile_layers.py:83-93``)
``ref_impala_step``    ``ImpalaLearner._train_step``                           ``rlego.vtrace_td_error_and_advantage``
                       (``agents/impala/learning.py:140-177``) incl. the       (shim = ``oracle/rlego_ref.py``; the V-trace
                       reference's ``functorch.vmap`` wrapper (``:15-18``),     recursion itself stays UNPINNED), everything
                       ``clip_grad_norm_`` and ``Adam.step``                    else is the reference's own op sequence
``ref_ppo_step``       ``PPOLearner._train_step``                              none -> **pinned by the reference itself**
                       (``agents/ppo/learning.py:130-143``) incl. clip + Adam
``ref_impala_loss``    ``losses.impala_loss`` (``losses.py:92-128``)           same shim
``ref_make_replay``    ``ImpalaActor._make_replay``                            none (``moolib.Batcher`` shim only hands the
                       (``agents/impala/learning.py:77-80``)                   stacked tensors back)
``ref_apex_replay``    ``ApexActor._async_make_replay``                        ``rlego.n_step_bootstrapped_returns`` (shim)
                       (``agents/dqn/learning.py:56-65``)                      -> pins slicing, discount, priorities
``ref_models``         ``models.models.AtariActorCritic`` over ``AtariBody``   none -> parameter counts / output shapes
                       and ``ImpalaCNNLarge`` (``models/common.py:64-126``)
=====================  =====================================================  ==========================================

The shims for ``rlmeta.utils.nested_utils`` restate three tree helpers ([EXT], rlmeta's documented behaviour):
``map_nested(fn, x)``, ``collate_nested(fn, [x_0..x_{B-1}])``, ``unbatch_nested(fn, x, B)``; they move tensors, they do
no arithmetic.  Nothing is written under ``/root/reference`` (``sys.dont_write_bytecode``).  The GPU box never runs
this file; it only reads the ``.npz`` outputs.
"""
import importlib
import os
import sys
import types

import numpy as np
import torch

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("IMPALA_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from impala_b200 import synth  # noqa: E402  (seeded synthetic rollouts only; no kernels)
from oracle import rlego_ref  # noqa: E402   (the rlego shim: the one unpinned piece)
from ref_cases import OutputsModel, impala_batch_list  # noqa: E402


# ------------------------------------------------------------------------------------------------ shims
def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _pkg(name, path=None, **attrs):
    m = _mod(name, **attrs)
    m.__path__ = [path] if path else []
    return m


def _map_nested(fn, nested):
    if isinstance(nested, (list, tuple)):
        return type(nested)(_map_nested(fn, x) for x in nested)
    if isinstance(nested, dict):
        return {k: _map_nested(fn, v) for k, v in nested.items()}
    return fn(nested)


def _collate_nested(fn, items):
    first = items[0]
    if isinstance(first, (list, tuple)):
        return type(first)(_collate_nested(fn, [it[i] for it in items]) for i in range(len(first)))
    if isinstance(first, dict):
        return {k: _collate_nested(fn, [it[k] for it in items]) for k in first}
    return fn(list(items))


def _unbatch_nested(fn, nested, batch_size):
    mapped = _map_nested(fn, nested)                           # fn sees the whole leaf (call sites unsqueeze dim 1), then dim 0 is unbound
    return [_map_nested(lambda x, i=i: x[i], mapped) for i in range(batch_size)]


class _Batcher:                                   # moolib.Batcher: only ever asked for what was stacked
    def __init__(self, *a, **k):
        self.data = None

    def get(self):
        return self.data

    def empty(self):
        return self.data is None


def install_shims():
    _mod("rlego",
         vtrace_td_error_and_advantage=rlego_ref.vtrace_td_error_and_advantage,
         lambda_returns=rlego_ref.lambda_returns,
         n_step_bootstrapped_returns=rlego_ref.n_step_bootstrapped_returns)
    _mod("moolib", Batcher=_Batcher)
    _mod("envs", EnvSpec=object, EnvFactory=object)
    _pkg("rlmeta")
    core = _pkg("rlmeta.core")
    core.remote = _mod("rlmeta.core.remote", Remote=type("Remote", (), {}),
                       remote_method=lambda **kw: (lambda f: f))
    _mod("rlmeta.core.model", ModelLike=object, RemotableModel=torch.nn.Module)
    _mod("rlmeta.core.replay_buffer", ReplayBufferLike=object)
    _mod("rlmeta.core.types", Action=tuple, TimeStep=tuple, NestedTensor=object)
    utils = _pkg("rlmeta.utils")
    utils.nested_utils = _mod("rlmeta.utils.nested_utils", map_nested=_map_nested,
                              collate_nested=_collate_nested, unbatch_nested=_unbatch_nested)
    # the reference's own packages: real files, but without running models/__init__.py (it star-imports the
    # rlmeta-dependent distributed models)
    _pkg("models", os.path.join(REF, "models"))
    _pkg("agents", os.path.join(REF, "agents"))
    _pkg("agents.impala", os.path.join(REF, "agents", "impala"))
    _pkg("agents.ppo", os.path.join(REF, "agents", "ppo"))
    _pkg("agents.dqn", os.path.join(REF, "agents", "dqn"))
    sys.path.insert(0, REF)
    sys.modules["models.distributed_models"] = _mod("models.distributed_models")   # only get_muesli (dead) uses it


def _np(d):
    return {k: (v.detach().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in d.items()}


def _save(name, **arrays):
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **_np(arrays))
    print("wrote", name)


# ------------------------------------------------------------------------------------------------ cases
def ref_ppo_loss(losses):
    keys = ("train/loss", "train/entropy", "train/td", "train/pg", "train/target", "train/kl", "train/ratio")
    for name, N, A, seed, ec, cc in (("ref_ppo_loss", 512, 15, 4100, 0.01, 0.1), ("ref_ppo_loss_a6", 96, 6, 4101, 0.02, 0.2)):
        d = synth.ppo_minibatch(N, A, seed)
        model = OutputsModel(d["logits"], d["v"])
        s = torch.zeros(N, 1)
        loss, m = losses.ppo_loss(model, (s, d["a"], d["v_target"], d["pi_ref"]), entropy_cost=ec, clip_coeff=cc)
        loss.backward()
        _save(name, **d, dlogits=model.logits.grad, dv=model.values.grad,
              metrics=np.array([m[k].item() for k in keys], dtype=np.float64),
              entropy_cost=np.float64(ec), clip_coeff=np.float64(cc))


def ref_quantile(ql):
    g = torch.Generator().manual_seed(4200)
    B, Ns, Nt = 64, 8, 8
    src = torch.randn(B, Ns, generator=g).requires_grad_(True)
    tau = torch.rand(B, Ns, generator=g)
    tgt = torch.randn(B, Nt, generator=g) * 2.0
    out = {}
    for tag, hp in (("h1", 1.0), ("h0", 0.0)):
        src.grad = None
        loss = ql.quantile_regression_loss(src, tau, tgt, hp)
        loss.sum().backward()
        out["loss_" + tag] = loss.detach().clone()
        out["dsrc_" + tag] = src.grad.clone()
    _save("ref_quantile", dist_src=src.detach(), tau_src=tau, dist_target=tgt, **out)


def ref_impala_step(learning):
    for name, B, T, A, seed in (("ref_impala_step", 32, 20, 15, 1000), ("ref_impala_step_small", 5, 7, 6, 4300)):
        d = synth.impala_rollout(B, T, A, seed)
        model = OutputsModel(d["logits"], d["values"])
        opt = torch.optim.Adam(model.parameters(), lr=1e-4, eps=1e-5)          # agents/impala/builder.py:43-44
        learner = learning.ImpalaLearner(model, None, opt, batch_size=B)       # max_grad_norm .5, entropy .01
        batch = impala_batch_list(d)
        captured = {}
        clip = torch.nn.utils.clip_grad_norm_

        def spy(params, max_norm, *a, **k):
            params = list(params)
            captured["dlogits"] = params[0].grad.clone()
            captured["dvalues"] = params[1].grad.clone()
            return clip(params, max_norm, *a, **k)

        torch.nn.utils.clip_grad_norm_ = spy
        try:
            m = learner._train_step(batch)
        finally:
            torch.nn.utils.clip_grad_norm_ = clip
        keys = ("train/loss", "train/entropy", "train/td", "train/pg", "train/kl", "train/ratio", "train/grad_norm")
        _save(name, **d, dlogits=captured["dlogits"], dvalues=captured["dvalues"],
              metrics=np.array([float(m[k]) for k in keys], dtype=np.float64),
              logits_after=model.logits.detach(), values_after=model.values.detach())


def ref_ppo_step(ppo_learning):
    """``PPOLearner._train_step`` (``agents/ppo/learning.py:130-143``): ``losses.ppo_loss`` -> backward -> clip -> Adam."""
    N, A = 256, 15                                                            # batch_size default, :81
    d = synth.ppo_minibatch(N, A, 4600)
    model = OutputsModel(d["logits"], d["v"])
    opt = torch.optim.Adam(model.parameters(), lr=1e-4, eps=1e-5)             # agents/ppo/builder.py
    learner = ppo_learning.PPOLearner(model, None, opt, batch_size=N)
    captured = {}
    clip = torch.nn.utils.clip_grad_norm_

    def spy(params, max_norm, *a, **k):
        params = list(params)
        captured["dlogits"], captured["dv"] = params[0].grad.clone(), params[1].grad.clone()
        return clip(params, max_norm, *a, **k)

    torch.nn.utils.clip_grad_norm_ = spy
    try:
        m = learner._train_step((torch.zeros(N, 1), d["a"], d["v_target"], d["pi_ref"]))
    finally:
        torch.nn.utils.clip_grad_norm_ = clip
    keys = ("train/loss", "train/entropy", "train/td", "train/pg", "train/target", "train/kl", "train/ratio",
            "train_step/grad_norm")
    assert set(m) == set(keys)
    _save("ref_ppo_step", **d, dlogits=captured["dlogits"], dv=captured["dv"],
          metrics=np.array([float(m[k]) for k in keys], dtype=np.float64),
          logits_after=model.logits.detach(), v_after=model.values.detach())


def ref_impala_loss(losses):
    T, A = 24, 15
    d = synth.impala_rollout(1, T, A, 4400)
    logits, values = d["logits"][0], d["values"][0]
    model = OutputsModel(logits, values)
    batch = (torch.zeros(T, 1), d["a"][0].unsqueeze(-1), d["r"][0].unsqueeze(-1), d["discount"][0].unsqueeze(-1),
             d["pi_ref"][0])
    loss, m = losses.impala_loss(batch, model)
    loss.backward()
    keys = ("train/loss", "train/td", "train/pg", "train/ratio")
    _save("ref_impala_loss", logits=logits, values=values, a=d["a"][0], r=d["r"][0], discount=d["discount"][0],
          pi_ref=d["pi_ref"][0], dlogits=model.logits.grad,
          dvalues=(model.values.grad if model.values.grad is not None else torch.zeros_like(values)),
          metrics=np.array([m[k].item() for k in keys], dtype=np.float64))


def ref_make_replay(learning):
    g = torch.Generator().manual_seed(4500)
    T, A = 20, 15
    s = torch.randint(0, 256, (T, 3, 8, 8), dtype=torch.uint8, generator=g)
    a = torch.randint(0, A, (T, 1), generator=g)
    r = torch.randn(T, 1, generator=g)
    done = torch.rand(T, 1, generator=g) < 0.2
    pi = torch.randn(T, A, generator=g)
    actor = learning.ImpalaActor(None, None, gamma=0.99, rollout_length=T)
    actor._trajectory.data = (s, a, r, done, pi)
    out = actor._make_replay()
    assert out[0] is s and out[1] is a and out[2] is r and out[4] is pi
    _save("ref_make_replay", done=done, discount=out[3], gamma=np.float64(0.99))


def ref_apex_replay(dqn_learning):
    """``ApexActor._async_make_replay`` (``agents/dqn/learning.py:56-65``): discount from done flags, 3-step bootstrapped
    targets on ``(r[:-1], discount[:-1], q_star[1:])``, priorities ``|target - q_pi[:-1]|``."""
    import asyncio
    g = torch.Generator().manual_seed(4700)
    T = 40
    s = torch.zeros(T, 1)
    a = torch.randint(0, 15, (T, 1), generator=g)
    r = torch.randn(T, 1, generator=g)
    done = torch.rand(T, 1, generator=g) < 0.1
    q_pi = torch.randn(T, 1, generator=g)
    q_star = q_pi + torch.rand(T, 1, generator=g)
    actor = dqn_learning.ApexActor(None, None, n_step=3, rollout_length=T, gamma=0.99)
    actor._trajectory.data = (s, a, r, done, q_pi, q_star)
    replay, priorities = asyncio.run(actor._async_make_replay())
    target = torch.cat([item[2] for item in replay])
    assert len(replay) == T - 1 and target.shape == (T - 1,)
    _save("ref_apex_replay", r=r.squeeze(1), done=done.squeeze(1), q_pi=q_pi.squeeze(1), q_star=q_star.squeeze(1),
          target=target, priorities=priorities, n_step=np.int64(3), gamma=np.float64(0.99))


def ref_models():
    mm = importlib.import_module("models.models")
    mc = importlib.import_module("models.common")
    rows = {}
    torch.manual_seed(0)
    ac = mm.AtariActorCritic((3, 64, 64), 15)
    rows["atari_ac_params"] = sum(p.numel() for p in ac.parameters())
    x = torch.rand(2, 3, 64, 64)
    p, v = ac(x)
    rows["atari_ac_logits_shape"], rows["atari_ac_value_shape"] = np.array(p.shape), np.array(v.shape)
    body = mc.ImpalaCNNLarge((3, 64, 64))
    rows["impala_cnn_params"] = sum(p.numel() for p in body.parameters())
    rows["impala_cnn_out"] = np.array(body(x).shape)
    rows["atari_body_params"] = sum(p.numel() for p in mc.AtariBody((3, 64, 64)).parameters())
    _save("ref_models", **rows)


def main():
    assert os.path.isdir(REF), REF + " missing: this generator only runs where the reference is mounted"
    install_shims()
    losses = importlib.import_module("losses")
    learning = importlib.import_module("agents.impala.learning")
    ql = importlib.import_module("models.quantile_layers")
    assert losses.__file__.startswith(REF) and learning.__file__.startswith(REF) and ql.__file__.startswith(REF)
    ref_ppo_loss(losses)
    ref_quantile(ql)
    ref_impala_step(learning)
    ref_ppo_step(importlib.import_module("agents.ppo.learning"))
    ref_impala_loss(losses)
    ref_make_replay(learning)
    ref_apex_replay(importlib.import_module("agents.dqn.learning"))
    ref_models()


if __name__ == "__main__":
    main()
