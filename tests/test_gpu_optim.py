"""GPU: K6 (fused global-norm clip + Adam) through the C ABI vs torch's own clip_grad_norm_ + Adam on the CPU
(the reference's arithmetic for agents/impala/learning.py:172-176)."""
import copy

import pytest
import torch

from gpu_util import assert_close
from impala_b200 import models, ops, optim

pytestmark = pytest.mark.gpu


def _ref_steps(p0, grads, max_norm, lr, betas, eps):
    p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([p], lr=lr, betas=betas, eps=eps)
    norms, clipped = [], []
    for g in grads:
        # clip_grad_norm_ with the norm taken in fp64: torch's fp32 CPU sum of >1e6 squares is itself ~2e-5 off
        norm = torch.linalg.vector_norm(g.double())
        coef = min(1.0, max_norm / (norm.item() + 1e-6)) if max_norm > 0 else 1.0
        p.grad = g * torch.tensor(coef, dtype=torch.float32)
        norms.append(norm.float())
        clipped.append(p.grad.clone())
        opt.step()
    st = opt.state[p]
    return p.detach(), st["exp_avg"], st["exp_avg_sq"], norms, clipped


@pytest.mark.parametrize("n", [1, 5, 1023, 4096, 630_352, 1_300_001])
@pytest.mark.parametrize("max_norm,scale", [(0.5, 1.0), (40.0, 1e-3), (0.0, 1.0)])
def test_clip_adam_matches_torch(cuda, n, max_norm, scale):
    g = torch.Generator().manual_seed(n)
    p0 = torch.randn(n, generator=g)
    grads = [scale * torch.randn(n, generator=g) * (1 + k) for k in range(4)]
    lr, betas, eps = 1e-4, (0.9, 0.999), 1e-5                 # agents/impala/builder.py:43-44
    want_p, want_m, want_v, want_norms, want_clipped = _ref_steps(p0, grads, max_norm, lr, betas, eps)
    p, m, v = p0.to(cuda), torch.zeros(n, device=cuda), torch.zeros(n, device=cuda)
    for k, gk in enumerate(grads):
        gd = gk.to(cuda)
        norm = ops.clip_adam_step(p, gd, m, v, k + 1, lr, betas[0], betas[1], eps, max_norm)
        assert_close(norm[0], want_norms[k], "norm", rtol=1e-5, atol=0)        # fp32 sum of n squares, different order
        assert_close(gd, want_clipped[k], "clipped grads", rtol=2e-5, atol=1e-9)
    assert_close(m, want_m, "exp_avg", rtol=2e-5, atol=1e-8)
    assert_close(v, want_v, "exp_avg_sq", rtol=4e-5, atol=1e-12)
    # a parameter moves ~lr per step; 4 steps of lr=1e-4 compared to 1e-7 absolute = 2.5e-4 of the motion
    assert_close(p, want_p, "params", rtol=1e-6, atol=1e-7)


def test_clip_adam_unaligned_and_deterministic(cuda):
    n = 70_001
    g = torch.Generator().manual_seed(1)
    base = [torch.randn(n + 1, generator=g).to(cuda) for _ in range(2)]
    outs = []
    for rep in range(2):
        bufs = [b.clone() for b in base] + [torch.zeros(n + 1, device=cuda), torch.zeros(n + 1, device=cuda)]
        p, gd, m, v = [b[1:] for b in bufs]                    # 4-byte aligned only: scalar path
        norm = ops.clip_adam_step(p, gd, m, v, 1, 1e-3, max_norm=0.5)
        outs.append((p.clone(), gd.clone(), m.clone(), v.clone(), norm.clone()))
    for a, b in zip(*outs):
        assert torch.equal(a, b)
    pa, ga, ma, va = [b.clone() for b in base] + [torch.zeros(n + 1, device=cuda), torch.zeros(n + 1, device=cuda)]
    norm = ops.clip_adam_step(pa[:n], ga[:n], ma[:n], va[:n], 1, 1e-3, max_norm=0.5)   # aligned: float4 path
    want_p, want_m, want_v, want_norms, _ = _ref_steps(base[0][:n].cpu(), [base[1][:n].cpu()], 0.5, 1e-3, (0.9, 0.999), 1e-8)
    assert_close(pa[:n], want_p, "params", rtol=1e-6, atol=1e-7)
    assert_close(norm[0], want_norms[0], "norm", rtol=1e-5, atol=0)


@pytest.mark.parametrize("body,channels_last", [("nature", False), ("impala", True)])
def test_fused_optimizer_class_matches_adam_on_model(cuda, body, channels_last):
    """Plain ``zero_grad() / backward() / step()`` use.  With channels_last convolution weights the flat buffers hold a
    weight in its physical order: step() must gather p.grad through the same strides."""
    torch.manual_seed(0)
    model = models.ActorCritic((3, 64, 64), 15, body=body, channels_last=channels_last).to(cuda)
    if channels_last:
        model = model.to(memory_format=torch.channels_last)
    ref = copy.deepcopy(model)
    opt = optim.FusedClipAdam(model.parameters(), lr=1e-4, eps=1e-5, max_grad_norm=0.5)
    ropt = torch.optim.Adam(ref.parameters(), lr=1e-4, eps=1e-5)
    obs = torch.randint(0, 256, (16, 3, 64, 64), device=cuda, dtype=torch.uint8)
    for _ in range(3):
        for mdl, o in ((model, opt), (ref, ropt)):
            o.zero_grad(set_to_none=True)
            pi, v = mdl(obs)
            (pi.logsumexp(-1).mean() + v.pow(2).mean()).backward()
            if o is ropt:
                torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.5)
            o.step()
    if body == "nature":
        for (n1, a), (_, b) in zip(model.named_parameters(), ref.named_parameters()):
            assert_close(a, b, n1, rtol=1e-5, atol=2e-7)
        return
    # IMPALA-CNN: the two copies' TF32 convolution gradients differ in their last bits from the second step on, and Adam's
    # m / sqrt(v) turns that into up to ~6 % of the motion for the smallest gradients (1.8e-5 seen on 6 of 2,304 elements).
    # Wrongly paired elements (a logical-order gather into the physical-order flat buffer) would move nearly EVERY element
    # of a 4-D weight by the whole motion, 3 steps x lr = 3e-4.
    for (n1, a), (_, b) in zip(model.named_parameters(), ref.named_parameters()):
        err = (a.detach() - b.detach()).abs()
        # no bound on the maximum: an element whose gradient is at noise level gets +-lr per step from either copy
        assert (err > 1e-5).float().mean().item() < 0.10, (n1, (err > 1e-5).float().mean().item())


def test_golden_next_rows(cuda):
    """Committed fixture (tests/golden/next_rows.npz, generated from torch's own Adam / the oracle)."""
    import os
    import numpy as np
    from conftest import GOLDEN
    from impala_b200 import vtrace
    z = np.load(os.path.join(GOLDEN, "next_rows.npz"))
    p = torch.from_numpy(z["adam_p0"]).to(cuda)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for k in range(3):
        norm = ops.clip_adam_step(p, torch.from_numpy(z["adam_grads"][k]).to(cuda), m, v, k + 1, 1e-4, eps=1e-5, max_norm=0.5)
        assert norm.item() == pytest.approx(z["adam_norms"][k], rel=1e-5)
    assert_close(p, torch.from_numpy(z["adam_p"]), "params", rtol=1e-6, atol=1e-7)
    assert_close(m, torch.from_numpy(z["adam_m"]), "exp_avg", rtol=2e-5, atol=1e-8)
    assert_close(v, torch.from_numpy(z["adam_v"]), "exp_avg_sq", rtol=4e-5, atol=1e-12)
    r, d, vv = (torch.from_numpy(z[k]).to(cuda) for k in ("r", "discount", "v"))
    assert np.array_equal(vtrace.n_step_bootstrapped_returns(r, d, vv, 3).cpu().numpy(), z["n_step3"])
    tree = torch.zeros(2048, dtype=torch.float64, device=cuda)
    ops.sumtree_update(tree, torch.arange(1000, device=cuda), torch.from_numpy(z["priorities"]).to(cuda))
    idx, prob = ops.sumtree_sample(tree, 1000, 123, 0, 0, 512)
    assert np.array_equal(idx.cpu().numpy(), z["st_idx"]) and np.array_equal(prob.cpu().numpy(), z["st_prob"])


def test_nan_gradient_poisons_the_update_like_torch(cuda):
    """torch.clamp propagates NaN through clip_grad_norm_'s coefficient (the reference: agents/impala/learning.py:172-173);
    fminf would swallow it: a non-finite norm must reach total_norm AND the parameters."""
    n = 5000
    p, m, v = torch.randn(n, device=cuda), torch.zeros(n, device=cuda), torch.zeros(n, device=cuda)
    g = torch.randn(n, device=cuda)
    g[123] = float("nan")
    norm = ops.clip_adam_step(p, g, m, v, 1, 1e-3, max_norm=0.5)
    assert torch.isnan(norm[0])
    assert torch.isnan(p).all(), "every parameter sees the NaN clip coefficient, as with torch.clamp"
    # the reference's arithmetic on the CPU
    q = torch.nn.Parameter(torch.randn(n))
    q.grad = g.cpu().clone()
    torch.nn.utils.clip_grad_norm_([q], 0.5)
    assert torch.isnan(q.grad).all()


def test_state_dict_round_trips_with_torch_adam(cuda):
    """FusedClipAdam.state_dict() is torch.optim.Adam's format (the base class saved an EMPTY state): a run
    checkpointed from the fused optimizer resumes in torch's Adam, and back, with the same moments and bias correction."""
    torch.manual_seed(3)
    shapes = [(16, 3, 3, 3), (16,), (40, 16)]
    pf = [torch.nn.Parameter(torch.randn(s, device=cuda)) for s in shapes]
    pt = [torch.nn.Parameter(p.detach().clone()) for p in pf]
    fused = optim.FusedClipAdam(pf, lr=1e-2, eps=1e-5)
    ref = torch.optim.Adam(pt, lr=1e-2, eps=1e-5)
    g = torch.Generator(device=cuda).manual_seed(1)

    def step_both(opt_a, pa, opt_b, pb):
        grads = [torch.randn(p.shape, generator=g, device=cuda) for p in pa]
        for p, q, gr in zip(pa, pb, grads):
            p.grad, q.grad = gr.clone(), gr.clone()
        opt_a.step()
        opt_b.step()

    for _ in range(3):
        step_both(fused, pf, ref, pt)
    sd = fused.state_dict()
    assert len(sd["state"]) == 3 and float(sd["state"][0]["step"]) == 3.0
    for i, p in enumerate(pt):          # same moments as torch's own state
        assert_close(sd["state"][i]["exp_avg"], ref.state[p]["exp_avg"], "exp_avg", rtol=1e-5, atol=1e-8)
    # fused -> torch: a fresh torch Adam resumed from the fused checkpoint continues like the fused one
    pt2 = [torch.nn.Parameter(p.detach().clone()) for p in pf]
    ref2 = torch.optim.Adam(pt2, lr=1e-2, eps=1e-5)
    ref2.load_state_dict(sd)
    # torch -> fused: a fresh fused optimizer resumed from torch's checkpoint
    pf2 = [torch.nn.Parameter(p.detach().clone()) for p in pt]
    fused2 = optim.FusedClipAdam(pf2, lr=1e-2, eps=1e-5)
    fused2.load_state_dict(ref.state_dict())
    assert fused2.step_count == 3
    step_both(fused, pf, ref2, pt2)
    for p, q in zip(pf, pt2):
        assert_close(p, q, "fused -> torch resume", rtol=1e-5, atol=1e-6)
    g = torch.Generator(device=cuda).manual_seed(2)
    grads = [torch.randn(p.shape, generator=g, device=cuda) for p in pt]
    for p, q, gr in zip(pt, pf2, grads):
        p.grad, q.grad = gr.clone(), gr.clone()
    ref.step()
    fused2.step()
    for p, q in zip(pt, pf2):
        assert_close(q, p, "torch -> fused resume", rtol=1e-5, atol=1e-6)


def test_device_resident_step_counter_matches_host_steps_and_replays_in_a_graph(cuda):
    """K6 with `device_counters`: the step number lives in device memory (a launch argument would be frozen into a CUDA
    graph).  Same parameters as the host-step path after every step, eagerly and as ONE captured launch replayed."""
    n = 70_000
    g = torch.Generator(device=cuda).manual_seed(11)
    p0 = torch.randn(n, generator=g, device=cuda)
    grads = [torch.randn(n, generator=g, device=cuda) * (1 + k) for k in range(6)]
    # host-step reference (the eager path tested against torch above)
    p, m, v = p0.clone(), torch.zeros(n, device=cuda), torch.zeros(n, device=cuda)
    want = []
    for k, gk in enumerate(grads):
        ops.clip_adam_step(p, gk.clone(), m, v, k + 1, 1e-3, eps=1e-5, max_norm=0.5)
        want.append(p.clone())
    # device counters, eager
    p, m, v = p0.clone(), torch.zeros(n, device=cuda), torch.zeros(n, device=cuda)
    ctr = torch.zeros(2, dtype=torch.int64, device=cuda)
    for k, gk in enumerate(grads[:3]):
        ops.clip_adam_step(p, gk.clone(), m, v, 0, 1e-3, eps=1e-5, max_norm=0.5, device_counters=ctr)
        assert torch.equal(p, want[k]), "step %d" % (k + 1)
    assert ctr.tolist() == [3, 0]
    # ... and one captured launch replayed three more times (the gradient buffer is refilled between replays)
    gbuf = torch.empty(n, device=cuda)
    side = torch.cuda.Stream(cuda)
    with torch.cuda.stream(side):
        ops.workspace(cuda)
        norm = torch.empty(1, device=cuda)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        ops.clip_adam_step(p, gbuf, m, v, 0, 1e-3, eps=1e-5, max_norm=0.5, total_norm=norm, device_counters=ctr)
    for k in range(3, 6):
        gbuf.copy_(grads[k])
        gr.replay()
        torch.cuda.synchronize()
        assert torch.equal(p, want[k]), "replayed step %d" % (k + 1)
    assert ctr.tolist() == [6, 0]
