"""Seeded synthetic rollouts of the shapes BASELINE.json names.  CPU tensors."""
import torch


def impala_rollout(B=32, T=20, A=15, seed=1000, ref_noise=0.3, p_done=0.02, gamma=0.99):
    """Batch-major ``[B,T,...]`` procgen-shaped learner inputs.  ``pi_ref = logits + noise`` so that
    rho straddles 1 and both clip branches are exercised."""
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn(B, T, A, generator=g)
    pi_ref = logits + ref_noise * torch.randn(B, T, A, generator=g)
    a = torch.randint(0, A, (B, T), generator=g)
    r = torch.randn(B, T, generator=g).clamp_(-10, 10)
    done = torch.rand(B, T, generator=g) < p_done
    discount = torch.logical_not(done) * gamma
    values = torch.randn(B, T, generator=g)
    return dict(logits=logits, values=values, a=a, r=r, discount=discount.float(), pi_ref=pi_ref)


def ppo_is_rollout(B=32, T=20, A=15, seed=3200, ref_noise=0.3, p_done=0.1):
    """Inputs of ``ppo_loss_is`` (reference ``losses.py:62``): an IMPALA-shaped rollout plus done flags and a
    bootstrap value per rollout."""
    d = impala_rollout(B, T, A, seed, ref_noise)
    g = torch.Generator().manual_seed(seed + 77)
    done = torch.rand(B, T, generator=g) < p_done
    return dict(logits=d["logits"], pi_ref=d["pi_ref"], a=d["a"], r=d["r"], values=d["values"], done=done,
                v_boot=torch.randn(B, generator=g))


def ppo_minibatch(N=2048, A=15, seed=3000, ref_noise=0.3):
    """Flat PPO minibatch: logits/pi_ref ``[N,A]``, v ``[N,1]``, a, v_target ``[N]``."""
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn(N, A, generator=g)
    pi_ref = logits + ref_noise * torch.randn(N, A, generator=g)
    a = torch.randint(0, A, (N,), generator=g)
    v = torch.randn(N, 1, generator=g)
    v_target = v.squeeze(-1) + torch.randn(N, generator=g)
    return dict(logits=logits, v=v, a=a, v_target=v_target, pi_ref=pi_ref)


def returns_rollout(B=64, T=256, seed=3100, p_done=0.02, gamma=0.99):
    """``r, done, v`` batch-major ``[B,T]`` for lambda-returns / GAE."""
    g = torch.Generator().manual_seed(seed)
    r = torch.randn(B, T, generator=g).clamp_(-10, 10)
    done = torch.rand(B, T, generator=g) < p_done
    v = torch.randn(B, T, generator=g)
    return dict(r=r, done=done, discount=(torch.logical_not(done) * gamma).float(), v=v)
