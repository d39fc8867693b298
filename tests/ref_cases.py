"""Shared by ``tests/golden/make_golden_from_reference.py`` (which RUNS the reference's files) and the tests that
replay the same cases through the oracle (CPU) and the CUDA drop-ins (GPU)."""
import numpy as np
import torch

IMPALA_STEP_KEYS = ("train/loss", "train/entropy", "train/td", "train/pg", "train/kl", "train/ratio", "train/grad_norm")
PPO_KEYS = ("train/loss", "train/entropy", "train/td", "train/pg", "train/target", "train/kl", "train/ratio")
PPO_STEP_KEYS = PPO_KEYS + ("train_step/grad_norm",)
IMPALA_LOSS_KEYS = ("train/loss", "train/td", "train/pg", "train/ratio")
IMPALA_FIELDS = ("logits", "values", "a", "r", "discount", "pi_ref")
PPO_FIELDS = ("logits", "v", "a", "v_target", "pi_ref")


class OutputsModel(torch.nn.Module):
    """A 'model' whose outputs ARE its parameters, so the learner's backward / clip / Adam act on (logits, values)
    directly and the fixture can hold the exact gradients and the parameters after the step."""

    def __init__(self, logits, values):
        super().__init__()
        self.logits = torch.nn.Parameter(logits.clone())
        self.values = torch.nn.Parameter(values.clone())

    def forward(self, s):
        n = s.shape[0]
        return self.logits.reshape(n, -1), self.values.reshape(n, 1)


def impala_batch_list(d):
    """The batch as ``replay.sample`` hands it to the reference learner (``agents/impala/learning.py:121,142``): a list
    of B rollouts, each ``[s, a [T,1], r [T,1], discount [T,1], pi_ref [T,A]]``.  ``s`` is a dummy (the model above only
    looks at its leading size)."""
    B, T = d["a"].shape
    return [[torch.zeros(T, 1, 1, device=d["a"].device), d["a"][b].unsqueeze(-1), d["r"][b].unsqueeze(-1),
             d["discount"][b].unsqueeze(-1), d["pi_ref"][b]] for b in range(B)]


def impala_loss_batch(z, device="cpu"):
    """The single-sequence batch of ``losses.impala_loss`` (``losses.py:92-98``): ``(s, a [N,1], r [N,1], discount [N,1], pi_ref)``."""
    t = {k: torch.from_numpy(np.asarray(z[k])).to(device) for k in ("a", "r", "discount", "pi_ref")}
    n = t["a"].shape[0]
    return (torch.zeros(n, 1, device=device), t["a"].unsqueeze(-1), t["r"].unsqueeze(-1), t["discount"].unsqueeze(-1),
            t["pi_ref"])
