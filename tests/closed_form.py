"""The closed-form forward+backward the K1 / K3 kernels implement, written out in torch (any dtype)
so that it can be checked against autograd of the oracle on the CPU before any GPU time is spent."""
import torch


def impala_closed_form(logits, values, a, r, disc, pi_ref, ec=0.01, lam=1.0, clip_rho=1.0, clip_pg=1.0,
                       detach_adv=False, detach_rho=False):
    B, T, A = logits.shape
    L = torch.log_softmax(logits, -1)
    Lr = torch.log_softmax(pi_ref, -1)
    p = L.exp()
    H = -(p * L).sum(-1)
    kl = (p * (L - Lr)).sum(-1)
    lp = L.gather(-1, a[..., None])[..., 0]
    rho = (lp - Lr.gather(-1, a[..., None])[..., 0]).exp()
    N1, N = B * (T - 1), B * T
    v, vn = values[:, :-1], values[:, 1:]
    rh, d, rr = rho[:, :-1], disc[:, :-1], r[:, :-1]
    td = rh.clamp(max=clip_rho) * (rr + d * vn - v)
    gc = d * (rh.clamp(max=1.0) * lam)
    e = torch.zeros_like(td)
    acc = torch.zeros_like(td[:, 0])
    for t in reversed(range(T - 1)):
        acc = td[:, t] + gc[:, t] * acc
        e[:, t] = acc
    tgt = e + v
    err = tgt - v
    qb = torch.cat([lam * tgt[:, 1:] + (1 - lam) * v[:, 1:], vn[:, -1:]], 1)
    q = rr + d * qb
    cpg = rh.clamp(max=clip_pg)
    qmv = q - v
    adv = cpg * qmv
    pg = (lp[:, :-1] * adv).sum() / N1
    vl = (err ** 2).sum() / N1
    ent = H.sum() / N
    loss = -pg + vl - ec * ent
    # backward
    dpg_dlp = adv.clone()
    if not detach_adv and not detach_rho:
        dpg_dlp = dpg_dlp + torch.where(rh <= clip_pg, lp[:, :-1] * rh * qmv, torch.zeros_like(rh))
    g_lp = torch.zeros_like(lp)
    g_lp[:, :-1] = -dpg_dlp / N1
    onehot = torch.nn.functional.one_hot(a, A).to(logits.dtype)
    dlogits = g_lp[..., None] * (onehot - p) + (ec / N) * p * (L + H[..., None])
    dv = torch.zeros_like(values)
    dv[:, :-1] += -2 * err / N1
    if not detach_adv:
        dv[:, :-1] += lp[:, :-1] * cpg / N1
        w = torch.full_like(d, 1 - lam)
        w[:, -1] = 1.0
        dv[:, 1:] += -lp[:, :-1] * cpg * d * w / N1
    metrics = dict(loss=loss, entropy=ent, td=vl, pg=pg, kl=kl.sum() / N, ratio=rho.sum() / N)
    return loss, metrics, dlogits, dv


def ppo_closed_form(logits, v, a, v_target, pi_ref, ec=0.01, clip=0.1):
    N, A = logits.shape
    v = v.reshape(-1)
    L = torch.log_softmax(logits, -1)
    Lr = torch.log_softmax(pi_ref, -1)
    p = L.exp()
    H = -(p * L).sum(-1)
    kl = (p * (L - Lr)).sum(-1)
    lp = L.gather(-1, a[..., None])[..., 0]
    ratio = (lp - Lr.gather(-1, a[..., None])[..., 0]).exp()
    adv = v_target - v
    pg1 = -(adv * ratio)
    pg2 = -ratio.clamp(1 - clip, 1 + clip) * adv
    pg = torch.maximum(pg1, pg2).sum() / N
    td = 0.5 * (adv ** 2).sum() / N
    ent = H.sum() / N
    loss = pg + td - ec * ent
    inside = (ratio >= 1 - clip) & (ratio <= 1 + clip)
    g_ratio = torch.where(inside | (pg1 > pg2), -adv, torch.zeros_like(adv))
    g_lp = g_ratio * ratio / N
    onehot = torch.nn.functional.one_hot(a, A).to(logits.dtype)
    dlogits = g_lp[..., None] * (onehot - p) + (ec / N) * p * (L + H[..., None])
    dv = -adv / N
    metrics = dict(loss=loss, entropy=ent, td=td, pg=pg, target=v_target.mean(), kl=kl.sum() / N, ratio=ratio.mean())
    return loss, metrics, dlogits, dv
