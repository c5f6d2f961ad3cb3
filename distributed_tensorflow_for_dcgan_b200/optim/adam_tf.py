"""TensorFlow-form Adam.

``tf.train.AdamOptimizer(lr, beta1)`` (image_train.py:109-112; beta2=0.999, eps=1e-8 by TF
default) computes ``lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t)`` and ``p -= lr_t * m / (sqrt(v) +
eps)`` -- eps is added to the *uncorrected* ``sqrt(v)``, which differs from ``torch.optim.Adam``
(SURVEY.md 2.4d).  This is the reference implementation the fused CUDA Adam is tested against.
"""
from __future__ import annotations

import math
from typing import Dict, Iterable, List

import torch


def adam_tf_step_(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, t: int,
                  lr: float = 2e-4, beta1: float = 0.5, beta2: float = 0.999, eps: float = 1e-8) -> None:
    """One in-place TF-form Adam update; ``t`` is the 1-based step count of this optimizer."""
    lr_t = lr * math.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t)
    m.mul_(beta1).add_(g, alpha=1.0 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
    p.sub_(lr_t * m / (v.sqrt() + eps))


class AdamTF:
    """Adam over a named parameter dict with TF slot naming (``<var>/Adam``, ``<var>/Adam_1``,
    ``beta1_power``, ``beta2_power``)."""

    def __init__(self, params: Dict[str, torch.Tensor], names: Iterable[str], lr: float = 2e-4,
                 beta1: float = 0.5, beta2: float = 0.999, eps: float = 1e-8):
        self.params = params
        self.names: List[str] = list(names)
        self.lr, self.beta1, self.beta2, self.eps = lr, beta1, beta2, eps
        self.t = 0
        self.m = {n: torch.zeros_like(params[n]) for n in self.names}
        self.v = {n: torch.zeros_like(params[n]) for n in self.names}

    @torch.no_grad()
    def step(self, grads: Dict[str, torch.Tensor]) -> None:
        self.t += 1
        for n in self.names:
            adam_tf_step_(self.params[n], grads[n], self.m[n], self.v[n], self.t, self.lr, self.beta1,
                          self.beta2, self.eps)

    def state_dict(self, prefix: str = "") -> Dict[str, torch.Tensor]:
        from ..models.manifest import adam_slot_names
        out = {}
        for n in self.names:
            a, b = adam_slot_names(n)
            out[a], out[b] = self.m[n], self.v[n]
        out[prefix + "beta1_power"] = torch.tensor(self.beta1 ** (self.t + 1))
        out[prefix + "beta2_power"] = torch.tensor(self.beta2 ** (self.t + 1))
        out[prefix + "adam_t"] = torch.tensor(self.t, dtype=torch.int64)
        return out

    def load_state_dict(self, sd: Dict[str, torch.Tensor], prefix: str = "") -> None:
        from ..models.manifest import adam_slot_names
        for n in self.names:
            a, b = adam_slot_names(n)
            if a in sd:
                self.m[n].copy_(sd[a])
                self.v[n].copy_(sd[b])
        if prefix + "adam_t" in sd:
            self.t = int(sd[prefix + "adam_t"])
