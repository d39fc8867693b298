"""Generates the committed golden vectors from the CPU oracle (run once here; the GPU box only reads
the .npz files).  The reference itself cannot be imported (rlego/rlmeta/moolib absent), so
these pin the ORACLE's outputs -- parity stays "unpinned" w.r.t. rlego (see oracle/__init__.py).

    python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from impala_b200 import synth  # noqa: E402
from oracle import losses_ref, replay_ref, rlego_ref  # noqa: E402


def impala_case(name, B, T, A, seed, **kw):
    d = synth.impala_rollout(B, T, A, seed)
    logits = d["logits"].clone().requires_grad_(True)
    values = d["values"].clone().requires_grad_(True)
    loss, m = losses_ref.impala_train_loss(logits, values, d["a"], d["r"], d["discount"], d["pi_ref"], **kw)
    loss.backward()
    adv, err, q = rlego_ref.batched(rlego_ref.vtrace_td_error_and_advantage)(
        d["values"][:, :-1], d["values"][:, 1:], d["r"][:, :-1], d["discount"][:, :-1],
        torch.ones_like(d["r"][:, :-1]))
    np.savez_compressed(os.path.join(HERE, name + ".npz"),
                        **{k: v.numpy() for k, v in d.items()},
                        dlogits=logits.grad.numpy(), dvalues=values.grad.numpy(),
                        metrics=np.array([m[k].item() for k in ("train/loss", "train/entropy", "train/td",
                                                                "train/pg", "train/kl", "train/ratio")], dtype=np.float64),
                        vtrace_rho1_adv=adv.numpy(), vtrace_rho1_err=err.numpy(), vtrace_rho1_q=q.numpy(),
                        kw=np.array(json.dumps(kw)))


def ppo_case(name, N, A, seed):
    d = synth.ppo_minibatch(N, A, seed)
    logits = d["logits"].clone().requires_grad_(True)
    v = d["v"].clone().requires_grad_(True)
    loss, m = losses_ref.ppo_loss_from_outputs(logits, v, d["a"], d["v_target"], d["pi_ref"], 0.01, 0.1)
    loss.backward()
    keys = ("train/loss", "train/entropy", "train/td", "train/pg", "train/target", "train/kl", "train/ratio")
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **{k: t.numpy() for k, t in d.items()},
                        dlogits=logits.grad.numpy(), dv=v.grad.numpy(),
                        metrics=np.array([m[k].item() for k in keys], dtype=np.float64))


def returns_case(name, B, T, seed):
    d = synth.returns_rollout(B, T, seed)
    lr = rlego_ref.batched(rlego_ref.lambda_returns, lambda_=0.95)(d["r"][:, :-1], d["discount"][:, :-1], d["v"][:, 1:])
    gae = rlego_ref.batched(lambda r, g, v: rlego_ref.truncated_generalized_advantage_estimation(r, g, 0.95, v))(
        d["r"][:, :-1], d["discount"][:, :-1], d["v"])
    np.savez_compressed(os.path.join(HERE, name + ".npz"), r=d["r"].numpy(), discount=d["discount"].numpy(),
                        v=d["v"].numpy(), lambda_returns=lr.numpy(), gae=gae.numpy())


def replay_case(name):
    idx = replay_ref.sample_uniform(123, 0, 0, 4096, 1_000_000)
    idx2 = replay_ref.sample_uniform(123, 3, 10_000_000_000, 257, 1000)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), idx_seed123_1M=idx, idx_stream3_big_offset=idx2)


def next_rows_case(name):
    """SURVEY sec 8f rows built in round 1: n-step returns (N4), sum-tree draws (N3), clip + Adam (N1)."""
    d = synth.returns_rollout(8, 24, 4100)
    ns = rlego_ref.batched_loop(rlego_ref.n_step_bootstrapped_returns, d["r"], d["discount"], d["v"], n=3)
    rng = np.random.default_rng(7)
    pri = rng.random(1000) ** 2 + 1e-3
    tree = replay_ref.SumTree(1000)
    tree.update(np.arange(1000), pri)
    idx, prob = tree.sample(123, 0, 0, 512, 1000)
    g = torch.Generator().manual_seed(11)
    p0, grads = torch.randn(4099, generator=g), [torch.randn(4099, generator=g) for _ in range(3)]
    p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([p], lr=1e-4, eps=1e-5)                    # reference agents/impala/builder.py:43-44
    norms = []
    for gk in grads:
        p.grad = gk.clone()
        norms.append(float(torch.nn.utils.clip_grad_norm_([p], 0.5)))  # reference agents/impala/learning.py:172
        opt.step()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), r=d["r"].numpy(), discount=d["discount"].numpy(),
                        v=d["v"].numpy(), n_step3=ns.numpy(), priorities=pri, st_idx=idx, st_prob=prob,
                        adam_p0=p0.numpy(), adam_grads=torch.stack(grads).numpy(), adam_p=p.detach().numpy(),
                        adam_m=opt.state[p]["exp_avg"].numpy(), adam_v=opt.state[p]["exp_avg_sq"].numpy(),
                        adam_norms=np.array(norms))


def ppo_is_case(name, B, T, A, seed):
    d = synth.ppo_is_rollout(B, T, A, seed)
    logits = d["logits"].clone().requires_grad_(True)
    v = d["values"].clone().requires_grad_(True)
    loss, m = losses_ref.ppo_is_loss_from_outputs(logits, v, d["v_boot"], d["a"], d["r"], d["done"], d["pi_ref"])
    loss.backward()
    keys = ("train/loss", "train/entropy", "train/td", "train/pg", "train/kl", "train/ratio")
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **{k: t.numpy() for k, t in d.items()},
                        dlogits=logits.grad.numpy(), dvalues=v.grad.numpy(),
                        metrics=np.array([m[k].item() for k in keys], dtype=np.float64))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "next_rows":       # add the new fixture without touching the others
        next_rows_case("next_rows")
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "ppo_is":
        ppo_is_case("ppo_is_small", 12, 20, 15, 3200)
        sys.exit(0)
    torch.manual_seed(0)
    impala_case("impala_c1", 32, 20, 15, 1000)                       # BASELINE config 1
    impala_case("impala_small_lambda", 5, 7, 6, 1001, lambda_=0.9, clip_rho_threshold=0.8,
                clip_pg_rho_threshold=1.2, entropy_coeff=0.05)
    ppo_case("ppo_small", 1000, 15, 3000)
    returns_case("returns_small", 16, 64, 3100)
    replay_case("replay_indices")
    next_rows_case("next_rows")
    ppo_is_case("ppo_is_small", 12, 20, 15, 3200)
    print("golden vectors written to", HERE)
