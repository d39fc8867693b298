"""GPU: K9 (fused IMPALA-CNN glue, csrc/cnn_glue.cu) against the torch ops the reference's body is written in
(``models/common.py:8-83``: ``nn.functional.relu`` / ``max_pool2d(kernel_size=3, stride=2, padding=1)`` / ``x + inputs``;
``models/models.py:73``: ``obs.float() / 255``).  Forward results are bit-exact; gradients differ only by the order
of <= 4 additions (pool) or of the bias-gradient reduction."""
import copy

import pytest
import torch
import torch.nn.functional as F

from gpu_util import assert_close
from impala_b200 import cnn_glue, models, ops

pytestmark = pytest.mark.gpu
CL = torch.channels_last


def _rand(shape, cuda, seed, levels=None):
    g = torch.Generator(device=cuda).manual_seed(seed)
    x = torch.randn(shape, device=cuda, generator=g)
    if levels:                                   # few distinct values -> many ties inside every window
        x = (x * levels).round() / levels
    return x.contiguous(memory_format=CL)


@pytest.mark.parametrize("shape", [(4, 3, 64, 64), (1, 3, 2, 2), (3, 3, 6, 10)])
def test_obs_conversion_bit_exact(cuda, shape):
    g = torch.Generator(device=cuda).manual_seed(1)
    obs = torch.randint(0, 256, shape, device=cuda, dtype=torch.uint8, generator=g)
    got = ops.obs_to_f32_nhwc(obs)
    want = obs.float() / 255.
    assert got.is_contiguous(memory_format=CL) and torch.equal(got, want)
    with pytest.raises(RuntimeError):
        ops.obs_to_f32_nhwc(obs[:, :2].contiguous())            # 2 planes: argument error from the C ABI


@pytest.mark.parametrize("shape,levels", [((5, 16, 64, 64), None), ((3, 32, 7, 9), 2), ((2, 8, 1, 1), None),
                                           ((2, 4, 2, 3), 1), ((7, 32, 16, 16), 3), ((1, 64, 5, 4), None)])
@pytest.mark.parametrize("with_bias", [True, False])
def test_maxpool_forward_and_backward_match_torch(cuda, shape, levels, with_bias):
    u = _rand(shape, cuda, 7, levels).requires_grad_(True)
    bias = torch.randn(shape[1], device=cuda) if with_bias else None
    ub = u + bias.view(1, -1, 1, 1) if with_bias else u
    want = F.max_pool2d(ub, kernel_size=3, stride=2, padding=1)
    p, r, code = ops.maxpool3x3s2_fwd(u.detach(), bias, with_relu=True)
    assert torch.equal(p, want) and torch.equal(r, F.relu(want))
    assert p.is_contiguous(memory_format=CL) and int(code.max()) <= 8
    gy = _rand(tuple(want.shape), cuda, 8)
    want.backward(gy)
    du = ops.maxpool3x3s2_bwd(gy, code, tuple(shape))
    assert_close(du, u.grad, "pool grad", rtol=1e-6, atol=1e-7)
    # every unit of gradient lands somewhere exactly once
    assert_close(du.sum(), gy.sum(), "mass", rtol=1e-4, atol=1e-3)


def test_maxpool_nan_rule(cuda):
    u = _rand((1, 4, 6, 6), cuda, 3)
    u[0, 1, 2, 2] = float("nan")
    p, _, _ = ops.maxpool3x3s2_fwd(u)
    want = F.max_pool2d(u, 3, 2, 1)
    assert torch.equal(torch.isnan(p), torch.isnan(want)) and torch.equal(p.nan_to_num(7.), want.nan_to_num(7.))


@pytest.mark.parametrize("shape", [(5, 16, 32, 32), (3, 32, 5, 7), (2, 4, 1, 1), (300, 32, 8, 8)])
def test_bias_act_and_relu_bwd_add_match_torch(cuda, shape):
    u, res = _rand(shape, cuda, 11, 2), _rand(shape, cuda, 12, 2)
    bias = torch.randn(shape[1], device=cuda)
    b4 = bias.view(1, -1, 1, 1)
    y, r = ops.bias_act_fwd(u, bias, res, want_y=True, want_r=True)
    assert torch.equal(y, (u + b4) + res) and torch.equal(r, F.relu((u + b4) + res))
    y2, r2 = ops.bias_act_fwd(u, bias, None, want_y=False, want_r=True)
    assert y2 is None and torch.equal(r2, F.relu(u + b4))
    dy, dr = _rand(shape, cuda, 13), _rand(shape, cuda, 14)
    g, db = ops.relu_bwd_add(dy, dr, r)
    want_g = dy + dr * (r > 0)
    assert torch.equal(g, want_g)
    assert_close(db, want_g.sum((0, 2, 3)), "dbias", rtol=1e-5, atol=1e-4)
    g2, db2 = ops.relu_bwd_add(None, dr, r)
    assert torch.equal(g2, dr * (r > 0))
    g3, db3 = ops.relu_bwd_add(dy, None, None)
    assert g3 is dy
    assert_close(db3, dy.sum((0, 2, 3)), "dbias colsum-only", rtol=1e-5, atol=1e-4)
    g4, db4 = ops.relu_bwd_add(dy, dr, r)
    assert torch.equal(db4, db)                                   # fixed-order sums: deterministic


def test_autograd_pieces_match_torch_block(cuda):
    """One ImpalaCNNBlock stage built from the fused pieces vs the same stage in eager torch (fp32 convs)."""
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        torch.manual_seed(3)
        stage = models._ImpalaStage(16, 32).to(cuda).to(memory_format=CL)
        x = _rand((6, 16, 20, 20), cuda, 5)
        xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
        want = stage(xa)
        u = F.conv2d(xb, stage.conv.weight, None, padding=1)
        p, r = cnn_glue.pool_bias(u, stage.conv.bias)
        for k, blk in enumerate((stage.res_block0, stage.res_block1)):
            r1 = cnn_glue.bias_relu(F.conv2d(r, blk.conv0.weight, None, padding=1), blk.conv0.bias)
            u1 = F.conv2d(r1, blk.conv1.weight, None, padding=1)
            if k == 0:
                p, r = cnn_glue.bias_add(u1, blk.conv1.bias, p, True, True)
            else:
                got = cnn_glue.bias_add(u1, blk.conv1.bias, p, True, False)
        assert_close(got, want, "stage output", rtol=1e-5, atol=1e-5)
        gy = _rand(tuple(want.shape), cuda, 6)
        grads_want = torch.autograd.grad(want, [xa] + list(stage.parameters()), gy)
        grads_got = torch.autograd.grad(got, [xb] + list(stage.parameters()), gy)
        for (n, _), a, b in zip([("x", None)] + list(stage.named_parameters()), grads_got, grads_want):
            assert_close(a, b, n, rtol=1e-4, atol=1e-5 * max(1., b.abs().max().item()), rel_to_max=1e-4)   # sums of ~10^4 terms
    finally:
        torch.backends.cudnn.allow_tf32 = tf32


@pytest.mark.parametrize("frame,n", [(64, 24), (128, 6), (84, 10), (50, 10)])
def test_fused_model_matches_eager_model(cuda, frame, n):
    """ActorCritic(body='impala', fused_glue=True) vs the eager channels_last model with the same parameters; 64x64 is
    procgen's frame (every convolution on K10/K11 or cuDNN's tcgen05 kernels), 128 / 84 / 50 put some or all layers outside
    the widths K10/K11 are built for -- those must route to cuDNN, with the K9 glue around them.  Both
    multiply their convolutions in TF32 (torch's default), so neither is "the" answer: an fp64 copy of the eager
    model is, and the fused model must be as close to it as the eager TF32 model is (within x8: the two carry
    different random rounding errors, here maxima over up to 9,216 elements), for outputs and for every parameter gradient."""
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True
    try:
        torch.manual_seed(0)
        eager = models.ActorCritic((3, frame, frame), 15, body="impala", channels_last=True).to(cuda).to(memory_format=CL)
        for p in eager.body.parameters():                           # biases are zero-initialised: make them matter
            if p.dim() == 1:
                torch.nn.init.normal_(p, std=0.1)
        fused = models.ActorCritic((3, frame, frame), 15, body="impala", channels_last=True, fused_glue=True).to(cuda).to(memory_format=CL)
        fused.load_state_dict(copy.deepcopy(eager.state_dict()))
        assert [n for n, _ in fused.named_parameters()] == [n for n, _ in eager.named_parameters()]
        exact = copy.deepcopy(eager).double()
        obs = torch.randint(0, 256, (n, 3, frame, frame), device=cuda, dtype=torch.uint8)
        outs = []
        for m in (exact, eager, fused):
            if m is exact:
                x = (obs.double() / 255.).contiguous(memory_format=CL)
                h = m.projection(m.body(x))
                pi, v = m.actor(h), m.critic(h)
            else:
                pi, v = m(obs)
            (pi.logsumexp(-1).mean() + (v - 1.).pow(2).mean() + 0.1 * pi.mean()).backward()
            outs.append((pi.detach().double(), v.detach().double()))

        def err(a, b):
            return (a.double() - b).abs().max().item() / max(b.abs().max().item(), 1e-30)

        for k, nm in enumerate(("logits", "values")):
            e_eager, e_fused = err(outs[1][k], outs[0][k]), err(outs[2][k], outs[0][k])
            assert e_fused <= 8 * e_eager + 2e-4, (nm, e_fused, e_eager)
        for (n, a), (_, b), (_, c) in zip(fused.named_parameters(), eager.named_parameters(), exact.named_parameters()):
            e_eager, e_fused = err(b.grad, c.grad), err(a.grad, c.grad)
            assert e_fused <= 8 * e_eager + 2e-4, (n, e_fused, e_eager)
    finally:
        torch.backends.cudnn.allow_tf32 = tf32


@pytest.mark.parametrize("N,Cin,Cout,H,W", [(6, 16, 16, 32, 32), (5, 3, 16, 64, 64), (4, 16, 32, 32, 32), (7, 32, 32, 16, 16),
                                            (9, 32, 32, 8, 8), (2, 32, 16, 12, 32), (1, 3, 32, 5, 8), (700, 32, 32, 8, 8)])
def test_conv3x3_wgrad_matches_cudnn_fp32(cuda, N, Cin, Cout, H, W):
    """K10 computes TF32 products with fp32 accumulation (what cuDNN does by default); the check is against cuDNN's
    fp32 (allow_tf32 = False) weight gradient at TF32 tolerance, plus exactness on TF32-representable inputs."""
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        x, dy = _rand((N, Cin, H, W), cuda, 21), _rand((N, Cout, H, W), cuda, 22)
        w = torch.randn(Cout, Cin, 3, 3, device=cuda).contiguous(memory_format=CL).requires_grad_(True)
        F.conv2d(x, w, None, padding=1).backward(dy)
        got = ops.conv3x3_wgrad(x, dy)
        assert got.shape == w.shape and got.stride() == w.grad.contiguous(memory_format=CL).stride()
        scale = w.grad.abs().max().item()
        assert (got - w.grad).abs().max().item() <= 2e-3 * scale, ((got - w.grad).abs().max().item(), scale)
        # inputs with <= 10 mantissa bits and small integer-like products: TF32 is exact, fp32 accumulation of
        # integers is exact -> bit-identical to the fp32 reference
        xi = (_rand((N, Cin, H, W), cuda, 23) * 4).round().clamp(-8, 8).contiguous(memory_format=CL)
        di = (_rand((N, Cout, H, W), cuda, 24) * 4).round().clamp(-8, 8).contiguous(memory_format=CL)
        w64 = w.detach().double().requires_grad_(True)                # fp64 reference: exact on integers whatever
        F.conv2d(xi.double(), w64, None, padding=1).backward(di.double())   # algorithm cuDNN picks
        assert torch.equal(ops.conv3x3_wgrad(xi, di), w64.grad.float())
        assert torch.equal(ops.conv3x3_wgrad(x, dy), got)            # deterministic
        with pytest.raises(RuntimeError):
            ops.conv3x3_wgrad(x[..., :4].contiguous(memory_format=CL), dy[..., :4].contiguous(memory_format=CL))   # W = 4
    finally:
        torch.backends.cudnn.allow_tf32 = tf32


@pytest.mark.parametrize("N,Cin,Cout,H,W", [(6, 16, 16, 32, 32), (5, 3, 16, 64, 64), (4, 16, 32, 32, 32), (7, 32, 32, 16, 16),
                                            (9, 32, 32, 8, 8), (2, 32, 16, 12, 32), (1, 3, 32, 5, 8), (300, 32, 32, 8, 8),
                                            (3, 16, 16, 3, 16), (2, 16, 32, 9, 64)])
def test_conv3x3_forward_and_dgrad_match_cudnn_fp32(cuda, N, Cin, Cout, H, W):
    """K11 (TF32 products, fp32 accumulation) vs cuDNN's fp32 convolution and its input gradient: TF32 tolerance on
    random data, bit-identical on TF32-exact small-integer data."""
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        w = (torch.randn(Cout, Cin, 3, 3, device=cuda) / (3 * Cin ** .5)).contiguous(memory_format=CL)
        x = _rand((N, Cin, H, W), cuda, 31).requires_grad_(True)
        dy = _rand((N, Cout, H, W), cuda, 32)
        want = F.conv2d(x, w, None, padding=1)
        got = ops.conv3x3_fwd(x.detach(), w)
        assert got.shape == want.shape and got.is_contiguous(memory_format=CL)
        assert (got - want).abs().max().item() <= 2e-3 * want.abs().max().item()
        if Cin != 3:
            want.backward(dy)
            dx = ops.conv3x3_dgrad(dy, w)
            assert dx.shape == x.shape and (dx - x.grad).abs().max().item() <= 2e-3 * x.grad.abs().max().item()
        xi = (_rand((N, Cin, H, W), cuda, 33) * 4).round().clamp(-8, 8).contiguous(memory_format=CL).requires_grad_(True)
        wi = (torch.randn(Cout, Cin, 3, 3, device=cuda) * 2).round().clamp(-4, 4).contiguous(memory_format=CL)
        di = (_rand((N, Cout, H, W), cuda, 34) * 4).round().clamp(-8, 8).contiguous(memory_format=CL)
        xi64 = xi.detach().double().requires_grad_(True)              # fp64 reference: exact on integers whatever
        wanti = F.conv2d(xi64, wi.double(), None, padding=1)           # algorithm cuDNN picks
        assert torch.equal(ops.conv3x3_fwd(xi.detach(), wi), wanti.float())
        if Cin != 3:
            wanti.backward(di.double())
            assert torch.equal(ops.conv3x3_dgrad(di, wi), xi64.grad.float())
    finally:
        torch.backends.cudnn.allow_tf32 = tf32


def test_cnn_kernels_against_cpu_torch_ops(cuda):
    """The same pieces against torch's CPU kernels (the reference's arithmetic on the host, fp32, no TF32 anywhere):
    pool / bias / ReLU / residual exactly, the convolution pieces at TF32 tolerance."""
    g = torch.Generator().manual_seed(5)
    N, Cin, Cout, H, W = 3, 16, 32, 16, 16
    x = torch.randn(N, Cin, H, W, generator=g)
    w = torch.randn(Cout, Cin, 3, 3, generator=g) / 12.
    b = torch.randn(Cout, generator=g)
    dy = torch.randn(N, Cout, H, W, generator=g)
    xc = x.clone().requires_grad_(True)
    wc = w.clone().requires_grad_(True)
    y_cpu = F.conv2d(xc, wc, None, padding=1)
    y_cpu.backward(dy)
    xg, wg, dyg = (t.to(cuda).contiguous(memory_format=CL) for t in (x, w, dy))
    for got, want in ((ops.conv3x3_fwd(xg, wg), y_cpu.detach()), (ops.conv3x3_dgrad(dyg, wg), xc.grad),
                      (ops.conv3x3_wgrad(xg, dyg), wc.grad)):
        assert (got.cpu() - want).abs().max().item() <= 2e-3 * want.abs().max().item()
    u = y_cpu.detach()
    pool_cpu = F.max_pool2d(u + b.view(1, -1, 1, 1), 3, 2, 1)
    p, r, code = ops.maxpool3x3s2_fwd(u.to(cuda).contiguous(memory_format=CL), b.to(cuda), with_relu=True)
    assert torch.equal(p.cpu(), pool_cpu) and torch.equal(r.cpu(), F.relu(pool_cpu))
    uc = u.clone().requires_grad_(True)
    gp = torch.randn(pool_cpu.shape, generator=g)
    F.max_pool2d(uc + b.view(1, -1, 1, 1), 3, 2, 1).backward(gp)
    du = ops.maxpool3x3s2_bwd(gp.to(cuda).contiguous(memory_format=CL), code, tuple(u.shape))
    assert_close(du, uc.grad, "pool grad vs CPU", rtol=1e-6, atol=1e-7)
    obs = torch.randint(0, 256, (2, 3, 8, 8), dtype=torch.uint8, generator=g)
    got = ops.obs_to_f32_nhwc(obs.to(cuda)).cpu()
    assert (got - obs.float() / 255.).abs().max().item() <= 6e-8      # CPU divides, CUDA multiplies by 1/255: 1 ulp


@pytest.mark.parametrize("N,C,Cout,H,W", [(5, 16, 16, 32, 32), (3, 16, 32, 9, 16), (4, 32, 32, 8, 8), (2, 3, 16, 6, 64)])
def test_conv3x3_fused_epilogue_exact_on_integers(cuda, N, C, Cout, H, W):
    """K11 with bias / residual / ReLU in its epilogue == conv2d(x, w, bias) (+ res) and its ReLU; bit-identical on
    TF32-exact integer data against an fp64 reference."""
    x = (_rand((N, C, H, W), cuda, 41) * 4).round().clamp(-8, 8).contiguous(memory_format=CL)
    w = (torch.randn(Cout, C, 3, 3, device=cuda) * 2).round().clamp(-4, 4).contiguous(memory_format=CL)
    b = (torch.randn(Cout, device=cuda) * 4).round()
    res = (_rand((N, Cout, H, W), cuda, 42) * 4).round().contiguous(memory_format=CL)
    conv = F.conv2d(x.double(), w.double(), b.double(), padding=1)
    y, r = ops.conv3x3_fwd_fused(x, w, b, res, want_y=True, want_r=True)
    assert torch.equal(y, (conv + res.double()).float()) and torch.equal(r, F.relu(conv + res.double()).float())
    y2, r2 = ops.conv3x3_fwd_fused(x, w, b, None, want_y=False, want_r=True)
    assert y2 is None and torch.equal(r2, F.relu(conv).float())
    y3, r3 = ops.conv3x3_fwd_fused(x, w, b, None, want_y=True, want_r=False)
    assert r3 is None and torch.equal(y3, conv.float())


@pytest.mark.parametrize("C", [16, 32])
def test_conv_bias_act_autograd_matches_torch_residual_block(cuda, C):
    """ResidualBlock (models/common.py:28-34) from the two fused pieces vs eager torch, forward and all gradients; the
    convolutions multiply in TF32 on both sides (K11 / cuDNN), so TF32 tolerance."""
    torch.manual_seed(7)
    blk = models._Residual(C).to(cuda).to(memory_format=CL)
    torch.nn.init.normal_(blk.conv0.bias, std=0.2)
    torch.nn.init.normal_(blk.conv1.bias, std=0.2)
    x = _rand((6, C, 16, 16), cuda, 43)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    want = blk(xa)
    r = F.relu(xb)
    r1 = cnn_glue.conv_bias_relu(r, blk.conv0.weight, blk.conv0.bias)
    got, got_r = cnn_glue.conv_bias_add(r1, blk.conv1.weight, blk.conv1.bias, xb, want_y=True, want_r=True)
    assert (got - want).abs().max().item() <= 3e-3 * want.abs().max().item()
    assert torch.equal(got_r, F.relu(got))
    gy, gr = _rand(tuple(want.shape), cuda, 44), _rand(tuple(want.shape), cuda, 45)
    grads_want = torch.autograd.grad((want * gy).sum() + (F.relu(want) * gr).sum(), [xa] + list(blk.parameters()))
    grads_got = torch.autograd.grad((got * gy).sum() + (got_r * gr).sum(), [xb] + list(blk.parameters()))
    for (n, _), a, b in zip([("x", None)] + list(blk.named_parameters()), grads_got, grads_want):
        assert (a - b).abs().max().item() <= 5e-3 * b.abs().max().item(), n
