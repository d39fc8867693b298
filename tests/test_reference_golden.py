"""CPU: pins the ORACLE against fixtures produced by running the reference's own files
(``tests/golden/make_golden_from_reference.py``; ``ref_*.npz``).  ``ref_ppo_loss*`` and ``ref_quantile`` involve no
third-party arithmetic, so for those rows the oracle is pinned by the reference itself; ``ref_impala_*`` pin
everything around the (still unpinned) rlego V-trace recursion: slices, means, signs, no-detach, clip, Adam."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import losses_ref
from ref_cases import (IMPALA_FIELDS, IMPALA_LOSS_KEYS, IMPALA_STEP_KEYS, PPO_FIELDS, PPO_KEYS, OutputsModel,
                       impala_loss_batch)


def load(name, fields):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: torch.from_numpy(z[k]) for k in fields}, z


@pytest.mark.parametrize("name", ["ref_ppo_loss", "ref_ppo_loss_a6"])
def test_oracle_ppo_loss_equals_reference_run(name):
    d, z = load(name, PPO_FIELDS)
    lg, v = d["logits"].clone().requires_grad_(True), d["v"].clone().requires_grad_(True)
    loss, m = losses_ref.ppo_loss_from_outputs(lg, v, d["a"], d["v_target"], d["pi_ref"], float(z["entropy_cost"]),
                                               float(z["clip_coeff"]))
    loss.backward()
    # same torch ops in the same order: identical, not just close
    assert np.array_equal(lg.grad.numpy(), z["dlogits"]) and np.array_equal(v.grad.numpy(), z["dv"])
    assert [m[k].item() for k in PPO_KEYS] == list(z["metrics"])
    # the reference signature (model, batch, ...) restated in the oracle (losses.py:131)
    model = OutputsModel(d["logits"], d["v"])
    loss2, _ = losses_ref.ppo_loss(model, (torch.zeros(len(d["a"]), 1), d["a"], d["v_target"], d["pi_ref"]),
                                   float(z["entropy_cost"]), float(z["clip_coeff"]))
    assert loss2.item() == z["metrics"][0]


def test_oracle_quantile_loss_equals_reference_run():
    d, z = load("ref_quantile", ("dist_src", "tau_src", "dist_target"))
    for tag, hp in (("h1", 1.0), ("h0", 0.0)):
        src = d["dist_src"].clone().requires_grad_(True)
        loss = losses_ref.quantile_regression_loss(src, d["tau_src"], d["dist_target"], hp)
        loss.sum().backward()
        assert np.array_equal(loss.detach().numpy(), z["loss_" + tag])
        assert np.array_equal(src.grad.numpy(), z["dsrc_" + tag])


@pytest.mark.parametrize("name", ["ref_impala_step", "ref_impala_step_small"])
def test_oracle_impala_train_loss_equals_reference_learner_step(name):
    """``ImpalaLearner._train_step`` (agents/impala/learning.py:140-177) ran with the oracle's rlego shim under the
    reference's own ``functorch.vmap``; the oracle's restatement of the loss must reproduce its metrics, the
    gradients it handed to ``clip_grad_norm_`` and the parameters Adam left behind."""
    d, z = load(name, IMPALA_FIELDS)
    lg, vv = d["logits"].clone().requires_grad_(True), d["values"].clone().requires_grad_(True)
    loss, m = losses_ref.impala_train_loss(lg, vv, d["a"], d["r"], d["discount"], d["pi_ref"])
    loss.backward()
    np.testing.assert_allclose(lg.grad.numpy(), z["dlogits"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(vv.grad.numpy(), z["dvalues"], rtol=1e-6, atol=1e-9)
    for i, k in enumerate(IMPALA_STEP_KEYS[:-1]):
        assert m[k].item() == pytest.approx(z["metrics"][i], rel=1e-6, abs=1e-9), k
    # clip (0.5) + Adam(lr 1e-4, eps 1e-5) as the reference builds them (agents/impala/builder.py:43-44)
    p = [d["logits"].clone().requires_grad_(True), d["values"].clone().requires_grad_(True)]
    p[0].grad, p[1].grad = torch.from_numpy(z["dlogits"]).clone(), torch.from_numpy(z["dvalues"]).clone()
    norm = torch.nn.utils.clip_grad_norm_(p, 0.5)
    torch.optim.Adam(p, lr=1e-4, eps=1e-5).step()
    assert norm.item() == pytest.approx(z["metrics"][-1], rel=1e-6)
    assert np.array_equal(p[0].detach().numpy(), z["logits_after"])
    assert np.array_equal(p[1].detach().numpy(), z["values_after"])


def test_oracle_ppo_loss_plus_clip_adam_equals_reference_ppo_learner_step():
    """``PPOLearner._train_step`` (agents/ppo/learning.py:130-143) is reference code end to end (no rlego)."""
    d, z = load("ref_ppo_step", PPO_FIELDS)
    p = [d["logits"].clone().requires_grad_(True), d["v"].clone().requires_grad_(True)]
    loss, m = losses_ref.ppo_loss_from_outputs(p[0], p[1], d["a"], d["v_target"], d["pi_ref"], 0.01, 0.1)
    loss.backward()
    assert np.array_equal(p[0].grad.numpy(), z["dlogits"]) and np.array_equal(p[1].grad.numpy(), z["dv"])
    assert [m[k].item() for k in PPO_KEYS] == list(z["metrics"][:-1])
    norm = torch.nn.utils.clip_grad_norm_(p, 0.5)
    torch.optim.Adam(p, lr=1e-4, eps=1e-5).step()
    assert norm.item() == z["metrics"][-1]
    assert np.array_equal(p[0].detach().numpy(), z["logits_after"]) and np.array_equal(p[1].detach().numpy(), z["v_after"])


def test_oracle_impala_loss_as_written_equals_reference_run():
    z = np.load(os.path.join(GOLDEN, "ref_impala_loss.npz"))
    model = OutputsModel(torch.from_numpy(z["logits"]), torch.from_numpy(z["values"]))
    loss, m = losses_ref.impala_loss(impala_loss_batch(z), model)
    loss.backward()
    assert [m[k].item() for k in IMPALA_LOSS_KEYS] == list(z["metrics"])
    assert np.array_equal(model.logits.grad.numpy(), z["dlogits"])
    assert np.array_equal(model.values.grad.numpy(), z["dvalues"])


def test_make_discount_and_model_shapes_equal_reference_run():
    from impala_b200 import models
    from impala_b200.agents.impala import make_discount
    z = np.load(os.path.join(GOLDEN, "ref_make_replay.npz"))
    got = make_discount(torch.from_numpy(z["done"]), float(z["gamma"]))
    assert np.array_equal(got.numpy(), z["discount"]) and got.dtype == torch.float32
    zm = np.load(os.path.join(GOLDEN, "ref_models.npz"))
    count = lambda m: sum(p.numel() for p in m.parameters())  # noqa: E731
    nature = models.ActorCritic((3, 64, 64), 15, body="nature")
    assert count(nature) == int(zm["atari_ac_params"])
    assert count(nature.body) == int(zm["atari_body_params"])
    impala = models.ActorCritic((3, 64, 64), 15, body="impala")
    assert count(impala.body) == int(zm["impala_cnn_params"])
    x = torch.rand(2, 3, 64, 64)
    p, v = nature(x)
    assert list(p.shape) == list(zm["atari_ac_logits_shape"]) and list(v.shape) == list(zm["atari_ac_value_shape"])
    assert list(impala.body(x).shape) == list(zm["impala_cnn_out"])


def test_oracle_n_step_targets_equal_reference_apex_actor():
    """``ApexActor._async_make_replay`` (agents/dqn/learning.py:56-65) ran here with the rlego shim: pins the call-site
    slicing ``(r[:-1], discount[:-1], q_star[1:])``, ``discount = logical_not(done) * gamma`` and the priorities."""
    from oracle import rlego_ref
    z = np.load(os.path.join(GOLDEN, "ref_apex_replay.npz"))
    r, done, q_pi, q_star = (torch.from_numpy(z[k]) for k in ("r", "done", "q_pi", "q_star"))
    discount = torch.logical_not(done) * float(z["gamma"])
    target = rlego_ref.n_step_bootstrapped_returns(r[:-1], discount[:-1], q_star[1:], n=int(z["n_step"]))
    assert np.array_equal(target.numpy(), z["target"])
    assert np.array_equal((target - q_pi[:-1]).abs().numpy(), z["priorities"])
