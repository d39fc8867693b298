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


def assert_grad_parity(got, want32, n, name, want64=None):
    """Gradient parity at the north_star tolerance, made non-vacuous: gradients of a mean over n samples are O(1/n), so
    they are compared UN-NORMALISED (x n) elementwise at rtol 1e-5 / atol 1e-6 against the fp32 oracle.  An element may
    miss that only when fp32 rounding explains it: a gradient element is a sum of O(10) terms that cancel, and torch's own
    fp32 result is then ~1e-5 from the fp64 truth on this scale (measured: profiles/r02_parity_errors.txt).  In that case
    the kernel must be as close to the fp64 oracle as torch's fp32 path is (within x3), and the fp64 oracle is required."""
    g = got.detach().double().cpu() * n
    w = want32.detach().double().cpu() * n
    assert g.shape == w.shape, (name, g.shape, w.shape)
    bad = (g - w).abs() > RTOL * w.abs() + ATOL
    if not bool(bad.any()):
        return
    assert want64 is not None, "%s: %d element(s) outside rtol 1e-5 / atol 1e-6 (x n) and no fp64 oracle given" % (name, int(bad.sum()))
    t = want64.detach().double().cpu() * n
    e_k, e_o = (g - t).abs().max().item(), (w - t).abs().max().item()
    # x3: two fp32 evaluations in different operation orders; their MAXIMA over thousands of elements differ by small
    # random factors (2.3x seen at T = 40 in the drawn-shape tests, 8.5e-6 vs 3.6e-6)
    assert e_k <= 3.0 * e_o + ATOL, "%s: kernel is %.3e from the fp64 oracle, torch fp32 is %.3e (x n scale), %d elements outside " \
        "the elementwise tolerance" % (name, e_k, e_o, int(bad.sum()))
    assert int(bad.sum()) <= max(16, g.numel() // 50), "%s: %d of %d elements outside the tolerance" % (name, int(bad.sum()), g.numel())


def to_cuda(d, device="cuda"):
    return {k: v.to(device) for k, v in d.items()}


def oracle_impala(d, dtype=torch.float32, **kw):
    from oracle import losses_ref
    lg = d["logits"].detach().to(dtype).clone().requires_grad_(True)
    vv = d["values"].detach().to(dtype).clone().requires_grad_(True)
    loss, m = losses_ref.impala_train_loss(lg, vv, d["a"], d["r"].to(dtype), d["discount"].to(dtype), d["pi_ref"].to(dtype), **kw)
    loss.backward()
    return loss.detach(), m, lg.grad, vv.grad


def npz_tensors(path, keys):
    z = np.load(path)
    return {k: torch.from_numpy(z[k]) for k in keys}, z
