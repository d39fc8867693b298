import numpy as np
import torch

# north_star tolerance for fp32 parity: rtol 1e-5 / atol 1e-6
RTOL, ATOL = 1e-5, 1e-6


def assert_close(got, want, name="", rtol=RTOL, atol=ATOL, rel_to_max=None):
    """north_star tolerance, plus (gradients are O(1/N), so atol alone would be vacuous) an error bound
    relative to the largest reference magnitude."""
    got = got.detach().float().cpu()
    want = want.detach().float().cpu()
    assert got.shape == want.shape, (name, got.shape, want.shape)
    torch.testing.assert_close(got, want, rtol=rtol, atol=atol, msg=lambda m: name + ": " + m)
    if rel_to_max is not None and want.numel():
        scale = want.abs().max().item()
        err = (got - want).abs().max().item()
        # 1e-10 floor: a reference that is exactly zero (A = 1: one action, no gradient) may come back as fp32
        # rounding residue of the factored form e_j (c1 x_j + k1) + g_lp
        assert err <= rel_to_max * scale + 1e-10, "%s: max err %.3e vs scale %.3e" % (name, err, scale)


def to_cuda(d, device="cuda"):
    return {k: v.to(device) for k, v in d.items()}


def oracle_impala(d, **kw):
    from oracle import losses_ref
    lg = d["logits"].clone().requires_grad_(True)
    vv = d["values"].clone().requires_grad_(True)
    loss, m = losses_ref.impala_train_loss(lg, vv, d["a"], d["r"], d["discount"], d["pi_ref"], **kw)
    loss.backward()
    return loss.detach(), m, lg.grad, vv.grad


def npz_tensors(path, keys):
    z = np.load(path)
    return {k: torch.from_numpy(z[k]) for k in keys}, z
