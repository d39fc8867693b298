"""Measured parity errors of the fused kernels (run on a GPU box; output committed as profiles/r02_parity_errors.txt).

For every case: the CUDA kernel (fp32), the oracle in fp32 (what the tests compare with) and the oracle in fp64
(ground truth).  Scalars are compared as they are; gradients are compared UN-NORMALISED (gradient x N, N = number of
samples the loss averages over), so that north_star's atol 1e-6 is not vacuous against values of order 1/N.
Columns: max |kernel - f64|, max |oracle32 - f64|, max |kernel - oracle32| and the number of elements violating
|kernel - oracle32| <= 1e-5 |oracle32| + 1e-6  (the north_star tolerance, elementwise)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from impala_b200 import ops, synth  # noqa: E402
from oracle import losses_ref, rlego_ref  # noqa: E402

RTOL, ATOL = 1e-5, 1e-6


def row(name, k, o32, o64, scale=1.0):
    k, o32, o64 = k.detach().double().cpu() * scale, o32.detach().double().cpu() * scale, o64.detach().double().cpu() * scale
    viol = int(((k - o32).abs() > RTOL * o32.abs() + ATOL).sum())
    print("  %-22s |ref| max %.3e | kernel-f64 %.3e | oracle32-f64 %.3e | kernel-oracle32 %.3e | violations %d / %d" % (
        name, o64.abs().max(), (k - o64).abs().max(), (o32 - o64).abs().max(), (k - o32).abs().max(), viol, k.numel()), flush=True)
    return viol


def impala_case(B, T, A, seed):
    d = synth.impala_rollout(B, T, A, seed=seed)
    g = {k: v.cuda() for k, v in d.items()}
    sc, dl, dv = ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"])
    torch.cuda.synchronize()
    res = {}
    for dt in (torch.float32, torch.float64):
        lg, vv = d["logits"].detach().to(dt).clone().requires_grad_(True), d["values"].detach().to(dt).clone().requires_grad_(True)
        loss, m = losses_ref.impala_train_loss(lg, vv, d["a"], d["r"].to(dt), d["discount"].to(dt), d["pi_ref"].to(dt))
        loss.backward()
        res[dt] = (m, lg.grad, vv.grad)
    print("K1 impala_loss  B=%d T=%d A=%d (gradients x N, N = B*T = %d)" % (B, T, A, B * T))
    v = 0
    for s, k in ((ops.S_LOSS, "train/loss"), (ops.S_ENTROPY, "train/entropy"), (ops.S_TD, "train/td"), (ops.S_PG, "train/pg"),
                 (ops.S_KL, "train/kl"), (ops.S_RATIO, "train/ratio")):
        v += row(k, sc[s], res[torch.float32][0][k], res[torch.float64][0][k])
    v += row("dlogits x N", dl, res[torch.float32][1], res[torch.float64][1], B * T)
    v += row("dvalues x N", dv, res[torch.float32][2], res[torch.float64][2], B * T)
    return v


def ppo_case(N, A, seed):
    p = synth.ppo_minibatch(N, A, seed=seed)
    g = {k: v.cuda() for k, v in p.items()}
    sc, dl, dv = ops.ppo_loss_fwdbwd(g["logits"], g["v"].reshape(-1), g["a"], g["v_target"], g["pi_ref"], 0.01, 0.1)
    torch.cuda.synchronize()
    res = {}
    for dt in (torch.float32, torch.float64):
        lg, vv = p["logits"].detach().to(dt).clone().requires_grad_(True), p["v"].detach().reshape(-1).to(dt).clone().requires_grad_(True)
        loss, m = losses_ref.ppo_loss_from_outputs(lg, vv, p["a"], p["v_target"].to(dt), p["pi_ref"].to(dt), 0.01, 0.1)
        loss.backward()
        res[dt] = (m, lg.grad, vv.grad)
    print("K3 ppo_loss  N=%d A=%d (gradients x N)" % (N, A))
    v = 0
    for s, k in ((ops.S_LOSS, "train/loss"), (ops.S_ENTROPY, "train/entropy"), (ops.S_TD, "train/td"), (ops.S_PG, "train/pg"),
                 (ops.S_KL, "train/kl"), (ops.S_RATIO, "train/ratio"), (ops.S_TARGET, "train/target")):
        v += row(k, sc[s], res[torch.float32][0][k], res[torch.float64][0][k])
    v += row("dlogits x N", dl, res[torch.float32][1], res[torch.float64][1], N)
    v += row("dv x N", dv, res[torch.float32][2], res[torch.float64][2], N)
    return v


def returns_case(B, T, seed):
    x = synth.returns_rollout(B, T + 1, seed=seed)
    r, dsc, vn = x["r"][:, :T], x["discount"][:, :T], x["v"][:, 1:]
    got = ops.lambda_returns(r.cuda(), dsc.cuda(), vn.cuda(), 0.95)
    torch.cuda.synchronize()
    fn = rlego_ref.batched(rlego_ref.lambda_returns, lambda_=0.95)
    o32, o64 = fn(r, dsc, vn), fn(r.double(), dsc.double(), vn.double())
    print("K2 lambda_returns  B=%d T=%d" % (B, T))
    return row("targets", got, o32, o64)


def main():
    print(__doc__)
    v = 0
    v += impala_case(32, 20, 15, 1000)
    v += impala_case(4096, 20, 15, 2000)
    v += impala_case(512, 20, 6, 2001)
    v += ppo_case(256 * 64, 15, 3000)
    v += ppo_case(256 * 1024, 15, 3001)
    v += returns_case(1024, 256, 4000)
    print("total violations of the north_star tolerance: %d" % v)


if __name__ == "__main__":
    main()
