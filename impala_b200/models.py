"""The actor-critic networks that produce the hot path's inputs (PyTorch/cuDNN, not hand kernels).

Only needed to close BASELINE config 5 (full learner step): same layer shapes and parameter counts
as the reference's ``AtariActorCritic`` (``models/models.py:61-76``) over either the Nature-CNN
``AtariBody`` (``models/common.py:108-126``, 344,496 params at A=15) or the IMPALA-CNN
``ImpalaCNNLarge`` (``models/common.py:64-83``, 630,352 params at A=15).
"""
import math
from typing import Tuple

import torch
from torch import nn
from torch.nn import functional as F

_TRUNC_STD = 0.87962566103423978  # std of a unit normal truncated to [-2, 2]


def _trunc_init(layer: nn.Module, scale: float = 1.0) -> nn.Module:
    """Variance-scaling truncated-normal init, zero bias (reference ``layer_init_truncated``,
    ``models/common.py:152-159``)."""
    w = layer.weight
    fan_in = w[0].numel()
    with torch.no_grad():
        nn.init.trunc_normal_(w, std=math.sqrt(scale / max(1, fan_in)) / _TRUNC_STD)
        layer.bias.zero_()
    return layer


class _Residual(nn.Module):
    def __init__(self, ch: int):
        super().__init__()
        self.conv0 = nn.Conv2d(ch, ch, 3, padding=1)
        self.conv1 = nn.Conv2d(ch, ch, 3, padding=1)

    def forward(self, x):
        y = self.conv0(F.relu(x))
        y = self.conv1(F.relu(y))
        return x + y


class _ImpalaStage(nn.Module):
    """conv3x3 -> maxpool3/2 -> two residual blocks (Espeholt et al. 2018, fig. 3 right)."""

    def __init__(self, cin: int, cout: int):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, 3, padding=1)
        self.res_block0 = _Residual(cout)
        self.res_block1 = _Residual(cout)

    def forward(self, x):
        x = F.max_pool2d(self.conv(x), kernel_size=3, stride=2, padding=1)
        return self.res_block1(self.res_block0(x))


class ImpalaCNN(nn.Module):
    """16-32-32 IMPALA-CNN; 64x64 input -> 32*8*8 = 2048 features.

    ``forward_fused`` computes the same function with the cuDNN convolutions called bias-free and everything between
    them (bias, ReLU, residual add, max-pool, and their backward passes incl. the bias gradients) in the K9 glue
    kernels (``cnn_glue.py``): CUDA float32 channels_last tensors only, same parameters, same state dict."""

    def __init__(self, obs_shape: Tuple[int, int, int]):
        super().__init__()
        c, h, w = obs_shape
        self.body = nn.Sequential(_ImpalaStage(c, 16), _ImpalaStage(16, 32), _ImpalaStage(32, 32),
                                  nn.ReLU(), nn.Flatten())
        for _ in range(3):
            h, w = (h + 1) // 2, (w + 1) // 2
        self.output_dim = 32 * h * w

    def forward(self, x):
        return self.body(x)

    def forward_fused(self, x):
        from . import cnn_glue as G
        stages = [m for m in self.body if isinstance(m, _ImpalaStage)]
        for si, st in enumerate(stages):
            u = G.conv3x3(x, st.conv.weight)
            p, r = G.pool_bias(u, st.conv.bias)                            # max_pool(conv + b), relu of it
            for bi, blk in enumerate((st.res_block0, st.res_block1)):
                r1 = G.conv_bias_relu(r, blk.conv0.weight, blk.conv0.bias)
                w1, b1 = blk.conv1.weight, blk.conv1.bias
                if bi == 0:                                                # block output + the next block's opening ReLU
                    p, r = G.conv_bias_add(r1, w1, b1, p, want_y=True, want_r=True)
                elif si + 1 < len(stages):                                 # feeds the next stage's convolution
                    x = G.conv_bias_add(r1, w1, b1, p, want_y=True, want_r=False)
                else:                                                      # the body's closing ReLU
                    x = G.conv_bias_add(r1, w1, b1, p, want_y=False, want_r=True)
        return x.flatten(1)


class NatureCNN(nn.Module):
    """conv8/4x32, conv4/2x64, conv3/1x64; 64x64 input -> 1024 features."""

    def __init__(self, obs_shape: Tuple[int, int, int]):
        super().__init__()
        c, h, w = obs_shape
        self.body = nn.Sequential(_trunc_init(nn.Conv2d(c, 32, 8, stride=4)), nn.ReLU(),
                                  _trunc_init(nn.Conv2d(32, 64, 4, stride=2)), nn.ReLU(),
                                  _trunc_init(nn.Conv2d(64, 64, 3, stride=1)), nn.ReLU(), nn.Flatten())
        for k, s in ((8, 4), (4, 2), (3, 1)):
            h, w = (h - k) // s + 1, (w - k) // s + 1
        self.output_dim = 64 * h * w

    def forward(self, x):
        return self.body(x)


class ActorCritic(nn.Module):
    """``x/255 -> body -> LayerNorm -> Linear(.,256) -> GELU -> {actor Linear(256,A), critic Linear(256,1)}``."""

    def __init__(self, obs_shape: Tuple[int, int, int], action_dim: int = 15, h_dim: int = 256,
                 body: str = "nature", channels_last: bool = False, fused_glue: bool = False):
        super().__init__()
        self.body = ImpalaCNN(obs_shape) if body == "impala" else NatureCNN(obs_shape)
        d = self.body.output_dim
        self.projection = nn.Sequential(nn.LayerNorm(d), _trunc_init(nn.Linear(d, h_dim)), nn.GELU())
        self.actor = _trunc_init(nn.Linear(h_dim, action_dim))
        self.critic = _trunc_init(nn.Linear(h_dim, 1))
        self.channels_last = channels_last
        if fused_glue and not (channels_last and body == "impala"):
            raise RuntimeError("fused_glue needs body='impala' and channels_last=True")
        self.fused_glue = fused_glue

    def forward(self, obs: torch.Tensor):
        if self.fused_glue:                                  # K9: CUDA only, raises otherwise (no silent fallback)
            from . import ops
            if obs.dtype == torch.uint8:
                x = ops.obs_to_f32_nhwc(obs.contiguous())    # obs.float() / 255 + NCHW -> NHWC in one pass
            else:
                x = (obs.float() / 255.).contiguous(memory_format=torch.channels_last)
            h = self.projection(self.body.forward_fused(x))
            return self.actor(h), self.critic(h)
        x = obs.float() / 255.
        if self.channels_last:
            x = x.contiguous(memory_format=torch.channels_last)
        h = self.projection(self.body(x))
        return self.actor(h), self.critic(h)

    @torch.inference_mode()
    def act(self, obs: torch.Tensor, deterministic_policy: torch.Tensor):
        """Reference ``AtariPPOModel.act`` (``models/distributed_models.py:21-32``): action, full logits, value."""
        logits, v = self.forward(obs.to(next(self.parameters()).device))
        from . import ops           # fused epilogue (N5): softmax -> categorical draw -> argmax select, one launch
        action = ops.act_epilogue(logits, deterministic_policy, self._act_seed, 0, self._act_draws)   # CUDA only
        self._act_draws += logits.shape[0]
        return action, logits, v

    _act_seed, _act_draws = 0, 0

    def reset_act_rng(self, seed: int) -> None:
        """Seed of the actor-side categorical draws (Philox stream; counter = rows drawn so far)."""
        self._act_seed, self._act_draws = int(seed), 0
