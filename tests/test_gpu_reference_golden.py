"""GPU: the CUDA drop-ins, called the way the reference calls its own functions, against fixtures produced by
RUNNING THE REFERENCE'S FILES (``tests/golden/make_golden_from_reference.py``): ``losses.ppo_loss`` and
``quantile_regression_loss`` are the reference's arithmetic end to end; ``ImpalaLearner._train_step`` /
``losses.impala_loss`` are the reference's code around the (unpinned) rlego V-trace shim."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from gpu_util import assert_close
from impala_b200 import losses, ops, optim
from impala_b200.agents.impala import ImpalaLearner
from impala_b200.agents.ppo import PPOLearner
from ref_cases import (IMPALA_FIELDS, IMPALA_LOSS_KEYS, IMPALA_STEP_KEYS, PPO_FIELDS, PPO_KEYS, PPO_STEP_KEYS, OutputsModel,
                       impala_batch_list, impala_loss_batch)

pytestmark = pytest.mark.gpu


def load(name, fields, device):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: torch.from_numpy(z[k]).to(device) for k in fields}, z


@pytest.mark.parametrize("name", ["ref_ppo_loss", "ref_ppo_loss_a6"])
def test_ppo_loss_dropin_equals_reference_run(cuda, name):
    d, z = load(name, PPO_FIELDS, cuda)
    model = OutputsModel(d["logits"], d["v"]).to(cuda)
    batch = (torch.zeros(len(d["a"]), 1, device=cuda), d["a"], d["v_target"], d["pi_ref"])
    loss, m = losses.ppo_loss(model, batch, entropy_cost=float(z["entropy_cost"]), clip_coeff=float(z["clip_coeff"]))
    loss.backward()
    assert set(m) == set(PPO_KEYS)
    for i, k in enumerate(PPO_KEYS):
        assert m[k].item() == pytest.approx(z["metrics"][i], rel=1e-5, abs=1e-6), k
    assert_close(model.logits.grad, torch.from_numpy(z["dlogits"]), "dlogits", rel_to_max=2e-6)
    assert_close(model.values.grad, torch.from_numpy(z["dv"]), "dv", rel_to_max=2e-6)


def test_quantile_loss_dropin_equals_reference_run(cuda):
    d, z = load("ref_quantile", ("dist_src", "tau_src", "dist_target"), cuda)
    for tag, hp in (("h1", 1.0), ("h0", 0.0)):
        src = d["dist_src"].clone().requires_grad_(True)
        loss = losses.quantile_regression_loss(src, d["tau_src"], d["dist_target"], hp)
        loss.sum().backward()
        assert_close(loss, torch.from_numpy(z["loss_" + tag]), "loss " + tag)
        assert_close(src.grad, torch.from_numpy(z["dsrc_" + tag]), "dsrc " + tag)


@pytest.mark.parametrize("name", ["ref_impala_step", "ref_impala_step_small"])
def test_k1_equals_reference_learner_gradients(cuda, name):
    d, z = load(name, IMPALA_FIELDS, cuda)
    sc, dl, dv = ops.impala_loss_fwdbwd(d["logits"], d["values"], d["a"], d["r"], d["discount"], d["pi_ref"])
    slots = (ops.S_LOSS, ops.S_ENTROPY, ops.S_TD, ops.S_PG, ops.S_KL, ops.S_RATIO)
    for i, s in enumerate(slots):
        assert sc[s].item() == pytest.approx(z["metrics"][i], rel=1e-5, abs=1e-6), IMPALA_STEP_KEYS[i]
    assert_close(dl, torch.from_numpy(z["dlogits"]), "dlogits", rel_to_max=2e-6)
    assert_close(dv, torch.from_numpy(z["dvalues"]), "dvalues", rel_to_max=2e-6)


@pytest.mark.parametrize("name", ["ref_impala_step", "ref_impala_step_small"])
@pytest.mark.parametrize("fused_opt", [True, False])
def test_impala_learner_dropin_equals_reference_train_step(cuda, name, fused_opt):
    """Same constructor arguments, same list-of-rollouts batch, same metric keys as the reference learner
    (agents/impala/learning.py:88-177); K1 + K6 (or torch's Adam) must leave the same metrics and parameters."""
    d, z = load(name, IMPALA_FIELDS, cuda)
    model = OutputsModel(d["logits"], d["values"]).to(cuda)
    if fused_opt:
        opt = optim.FusedClipAdam(model.parameters(), lr=1e-4, eps=1e-5)
    else:
        opt = torch.optim.Adam(model.parameters(), lr=1e-4, eps=1e-5)
    learner = ImpalaLearner(model, None, opt, batch_size=d["a"].shape[0])
    m = learner._train_step(impala_batch_list(d))
    assert set(m) == set(IMPALA_STEP_KEYS)
    for i, k in enumerate(IMPALA_STEP_KEYS):
        assert float(m[k]) == pytest.approx(z["metrics"][i], rel=1e-5, abs=1e-6), k
    assert_close(model.logits, torch.from_numpy(z["logits_after"]), "logits after Adam", rtol=1e-6, atol=2e-7)
    assert_close(model.values, torch.from_numpy(z["values_after"]), "values after Adam", rtol=1e-6, atol=2e-7)
    # the step really moved the parameters (lr 1e-4 * g/(|g|+eps))
    assert (model.logits.detach().cpu() - d["logits"].cpu()).abs().max().item() > 1e-5


@pytest.mark.parametrize("fused_opt", [True, False])
def test_ppo_learner_dropin_equals_reference_train_step(cuda, fused_opt):
    """agents/ppo/learning.py:130-143 ran here unmodified; K3 + K6 (or torch's Adam) must leave the same metrics,
    gradient norm and parameters."""
    d, z = load("ref_ppo_step", PPO_FIELDS, cuda)
    model = OutputsModel(d["logits"], d["v"]).to(cuda)
    if fused_opt:
        opt = optim.FusedClipAdam(model.parameters(), lr=1e-4, eps=1e-5)
    else:
        opt = torch.optim.Adam(model.parameters(), lr=1e-4, eps=1e-5)
    learner = PPOLearner(model, None, opt, batch_size=len(d["a"]))
    m = learner._train_step((torch.zeros(len(d["a"]), 1, device=cuda), d["a"], d["v_target"], d["pi_ref"]))
    assert set(m) == set(PPO_STEP_KEYS)
    for i, k in enumerate(PPO_STEP_KEYS):
        assert float(m[k]) == pytest.approx(z["metrics"][i], rel=1e-5, abs=1e-6), k
    assert_close(model.logits, torch.from_numpy(z["logits_after"]), "logits after Adam", rtol=1e-6, atol=2e-7)
    assert_close(model.values, torch.from_numpy(z["v_after"]), "v after Adam", rtol=1e-6, atol=2e-7)


def test_impala_loss_as_written_equals_reference_run(cuda):
    z = np.load(os.path.join(GOLDEN, "ref_impala_loss.npz"))
    model = OutputsModel(torch.from_numpy(z["logits"]), torch.from_numpy(z["values"])).to(cuda)
    loss, m = losses.impala_loss(impala_loss_batch(z, cuda), model, as_written=True)
    loss.backward()
    assert set(m) == set(IMPALA_LOSS_KEYS)
    for i, k in enumerate(IMPALA_LOSS_KEYS):
        assert m[k].item() == pytest.approx(z["metrics"][i], rel=1e-5, abs=1e-6), k
    assert_close(model.logits.grad, torch.from_numpy(z["dlogits"]), "dlogits", rel_to_max=2e-6)
    assert_close(model.values.grad, torch.from_numpy(z["dvalues"]), "dvalues", rel_to_max=2e-6)


def test_n_step_returns_equal_reference_apex_actor(cuda):
    """N4 kernel on the exact tensors the reference's ApexActor hands to rlego (agents/dqn/learning.py:56-65): bit-exact."""
    from impala_b200 import vtrace
    from impala_b200.agents.impala import make_discount
    z = np.load(os.path.join(GOLDEN, "ref_apex_replay.npz"))
    r, done, q_pi, q_star = (torch.from_numpy(z[k]).to(cuda) for k in ("r", "done", "q_pi", "q_star"))
    discount = make_discount(done, float(z["gamma"]))
    target = vtrace.n_step_bootstrapped_returns(r[:-1], discount[:-1], q_star[1:], n=int(z["n_step"]))
    assert torch.equal(target.cpu(), torch.from_numpy(z["target"]))
    assert torch.equal((target - q_pi[:-1]).abs().cpu(), torch.from_numpy(z["priorities"]))
