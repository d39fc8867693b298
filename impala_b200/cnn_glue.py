"""Autograd pieces of the fused IMPALA-CNN glue (K9, ``csrc/cnn_glue.cu``).

The reference's body (``models/common.py:8-83``) is ``conv -> max_pool -> 2 x [relu -> conv -> relu -> conv -> +x]``
per stage; the convolutions stay in cuDNN (bias-free here), and everything between them -- bias adds, ReLUs, residual
adds, the pools, and in the backward pass the ReLU masks, gradient joins and bias gradients -- runs in the kernels
below, one pass each, instead of ~10 eager elementwise / pooling launches per block.
"""
import torch

from . import ops

_CL = torch.channels_last


def _cl(g):
    return None if g is None else g.contiguous(memory_format=_CL)


class PoolBias(torch.autograd.Function):
    """``p = max_pool2d(u + bias, 3, 2, 1)`` and ``r = relu(p)`` (``ImpalaCNNBlock.forward``, ``models/common.py:53-55``,
    plus the ReLU that opens the first residual block, ``:30``)."""

    @staticmethod
    def forward(ctx, u, bias):
        p, r, code = ops.maxpool3x3s2_fwd(u, bias, with_relu=True)
        ctx.save_for_backward(code, r)
        ctx.in_shape = tuple(u.shape)
        return p, r

    @staticmethod
    def backward(ctx, dp, dr):
        code, r = ctx.saved_tensors
        g, db = ops.relu_bwd_add(_cl(dp), _cl(dr), r)           # gradient w.r.t. p, joined; its channel sums = d bias
        return ops.maxpool3x3s2_bwd(g, code, ctx.in_shape), db


class BiasAct(torch.autograd.Function):
    """``v = u + bias (+ res)``; outputs ``v`` and/or ``relu(v)``:
    ``relu(conv0(.))`` (``models/common.py:31-32``) is ``(want_y, want_r) = (False, True)`` without ``res``;
    ``conv1(.) + inputs`` (``:33-34``) has ``res`` and also emits the ReLU that opens the next block / closes the body."""

    @staticmethod
    def forward(ctx, u, bias, res, want_y, want_r):
        y, r = ops.bias_act_fwd(u, bias, res, want_y, want_r)
        ctx.save_for_backward(r) if want_r else ctx.save_for_backward()
        ctx.flags = (res is not None, want_y, want_r)
        return tuple(t for t in (y, r) if t is not None)

    @staticmethod
    def backward(ctx, *grads):
        has_res, want_y, want_r = ctx.flags
        grads = list(grads)
        dy = _cl(grads.pop(0)) if want_y else None
        dr = _cl(grads.pop(0)) if want_r else None
        r = ctx.saved_tensors[0] if want_r else None
        g, db = ops.relu_bwd_add(dy, dr, r)
        return g, db, (g if has_res else None), None, None


class Conv3x3(torch.autograd.Function):
    """Bias-free ``conv2d(x, weight, padding=1)`` with a 3x3 kernel (``models/common.py:12-26,41-45``): forward and input
    gradient in K11 (32 -> 32 channels: cuDNN, see ``_k11_wins``), weight gradient in K10 (``csrc/conv3x3.cu``) -- TF32 products with fp32 accumulation, torch's
    default arithmetic for these layers.  cuDNN's kernels for 3/16/32 channels are 3-10x off the HBM roofline."""

    @staticmethod
    def forward(ctx, x, weight):
        ctx.save_for_backward(x, weight)
        if _k11_wins(weight):
            return ops.conv3x3_fwd(x, weight)
        return torch.nn.functional.conv2d(x, weight, None, padding=1)

    @staticmethod
    def backward(ctx, gy):
        x, weight = ctx.saved_tensors
        gy = _cl(gy)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            if _k11_wins(weight):
                dx = ops.conv3x3_dgrad(gy, weight)
            else:
                dx = torch.ops.aten.convolution_backward(gy, x, weight, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1,
                                                         (True, False, False))[0]
        if ctx.needs_input_grad[1]:
            dw = ops.conv3x3_wgrad(x, gy)
        return dx, dw


def _k11_wins(weight):
    """Measured at 5,120 frames (``scripts/bench_conv.py``): K11 beats cuDNN 1.8-4.9x wherever a channel count is below
    32 (cuDNN falls back to sm80 kernels there); for 32 -> 32 cuDNN's tcgen05 kernels are 1.5x faster than K11's
    mma.sync path, so those layers keep cuDNN for the forward and the input gradient."""
    return not (weight.shape[0] == 32 and weight.shape[1] == 32)


def conv3x3(x, weight):
    if tuple(weight.shape[2:]) != (3, 3) or not ops.conv3x3_wgrad_supported(x.shape, weight.shape[0]):
        return torch.nn.functional.conv2d(x, weight, None, padding=1)      # shapes outside K10/K11: plain cuDNN autograd
    return Conv3x3.apply(x, weight)


class ConvBiasAct(torch.autograd.Function):
    """One convolution of a residual block with its elementwise tail (``models/common.py:30-34``):
    ``v = conv3x3(x, weight) + bias (+ res)``, outputs ``v`` and/or ``relu(v)``.  Where K11 runs the convolution the
    tail is its epilogue (the raw convolution output never reaches HBM); for 32 -> 32 channels cuDNN convolves and the
    K9 ``bias_act`` kernel does the tail.  Backward: ``g = dv + d relu * (relu > 0)`` with the bias gradient in the same
    pass (K9), then K11 / cuDNN input gradient and K10 weight gradient of ``g``; ``res`` receives ``g`` itself."""

    @staticmethod
    def forward(ctx, x, weight, bias, res, want_y, want_r):
        if _k11_wins(weight) and ops.conv3x3_wgrad_supported(x.shape, weight.shape[0]):
            y, r = ops.conv3x3_fwd_fused(x, weight, bias, res, want_y, want_r)
        else:
            y, r = ops.bias_act_fwd(torch.nn.functional.conv2d(x, weight, None, padding=1), bias, res, want_y, want_r)
        ctx.save_for_backward(*((x, weight, r) if want_r else (x, weight)))
        ctx.flags = (res is not None, want_y, want_r)
        return tuple(t for t in (y, r) if t is not None)

    @staticmethod
    def backward(ctx, *grads):
        has_res, want_y, want_r = ctx.flags
        x, weight = ctx.saved_tensors[:2]
        r = ctx.saved_tensors[2] if want_r else None
        grads = list(grads)
        dy = _cl(grads.pop(0)) if want_y else None
        dr = _cl(grads.pop(0)) if want_r else None
        g, db = ops.relu_bwd_add(dy, dr, r if dr is not None else None)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            if _k11_wins(weight) and ops.conv3x3_wgrad_supported(x.shape, weight.shape[0]):
                dx = ops.conv3x3_dgrad(g, weight)
            else:
                dx = torch.ops.aten.convolution_backward(g, x, weight, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1,
                                                         (True, False, False))[0]
        if ctx.needs_input_grad[1]:
            if ops.conv3x3_wgrad_supported(x.shape, weight.shape[0]):
                dw = ops.conv3x3_wgrad(x, g)
            else:
                dw = torch.ops.aten.convolution_backward(g, x, weight, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1,
                                                         (False, True, False))[1]
        return dx, dw, db, (g if has_res else None), None, None


def conv_bias_relu(x, weight, bias):
    """``relu(conv3x3(x) + bias)`` (``models/common.py:31-32``)."""
    return ConvBiasAct.apply(x, weight, bias, None, False, True)[0]


def conv_bias_add(x, weight, bias, res, want_y=True, want_r=False):
    """``res + (conv3x3(x) + bias)`` and/or its ReLU (``models/common.py:33-34`` + the next block's ``:30``)."""
    out = ConvBiasAct.apply(x, weight, bias, res, want_y, want_r)
    return out if len(out) > 1 else out[0]


def pool_bias(u, bias):
    return PoolBias.apply(u, bias)


def bias_relu(u, bias):
    return BiasAct.apply(u, bias, None, False, True)[0]


def bias_add(u, bias, res, want_y=True, want_r=False):
    out = BiasAct.apply(u, bias, res, want_y, want_r)
    return out if len(out) > 1 else out[0]
