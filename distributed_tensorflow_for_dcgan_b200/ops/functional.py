"""Plain-PyTorch statements of every op the reference graph uses, with TensorFlow geometry.

These are (a) the correctness oracle for the sm_100a kernels and (b) the CPU / gloo plumbing path
(BASELINE.json config 1).  All activations are NHWC and all weights are in TensorFlow layouts, so
the same parameter tensors feed both this path and the CUDA engine.

Geometry notes: ``padding='SAME'`` with k=5, s=2 on an even extent pads one
pixel before and two after; ``conv2d_transpose`` is the data-gradient of that conv, i.e. the
full transposed conv cropped by one at the top/left and two at the bottom/right.  Neither equals
PyTorch's symmetric ``padding=2`` / ``output_padding=1``.
"""
from __future__ import annotations

import math
from typing import Tuple

import torch
import torch.nn.functional as F


def same_padding(in_size: int, k: int = 5, s: int = 2) -> Tuple[int, int, int]:
    """TensorFlow SAME: returns (out_size, pad_before, pad_after)."""
    out = -(-in_size // s)
    total = max((out - 1) * s + k - in_size, 0)
    before = total // 2
    return out, before, total - before


def conv2d_same(x: torch.Tensor, w_hwio: torch.Tensor, bias: torch.Tensor | None = None,
                stride: int = 2) -> torch.Tensor:
    """``tf.nn.conv2d(x, w, [1,s,s,1], 'SAME') + bias`` (distriubted_model.py:183-185). NHWC in/out."""
    kh, kw = w_hwio.shape[0], w_hwio.shape[1]
    _, pt, pb = same_padding(x.shape[1], kh, stride)
    _, pl, pr = same_padding(x.shape[2], kw, stride)
    xn = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    y = F.conv2d(xn, w_hwio.permute(3, 2, 0, 1), bias, stride=stride)
    return y.permute(0, 2, 3, 1)


def deconv2d(x: torch.Tensor, w_hwoi: torch.Tensor, bias: torch.Tensor | None = None,
             stride: int = 2) -> torch.Tensor:
    """``tf.nn.conv2d_transpose(x, w, output_shape=2x, [1,s,s,1]) + bias``
    (distriubted_model.py:200-208).  ``w`` is ``[kh, kw, Cout, Cin]``.  NHWC in/out."""
    kh, kw = w_hwoi.shape[0], w_hwoi.shape[1]
    oh, ow = x.shape[1] * stride, x.shape[2] * stride
    _, pt, _ = same_padding(oh, kh, stride)
    _, pl, _ = same_padding(ow, kw, stride)
    full = F.conv_transpose2d(x.permute(0, 3, 1, 2), w_hwoi.permute(3, 2, 0, 1), None, stride=stride)
    y = full[:, :, pt:pt + oh, pl:pl + ow]
    if bias is not None:
        y = y + bias.view(1, -1, 1, 1)
    return y.permute(0, 2, 3, 1)


def linear(x: torch.Tensor, matrix: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    """``tf.matmul(x, Matrix) + bias`` (distriubted_model.py:171-173)."""
    return x @ matrix + bias


def lrelu(x: torch.Tensor, leak: float = 0.2) -> torch.Tensor:
    """``tf.maximum(x, leak*x)`` (distriubted_model.py:156-157)."""
    return torch.maximum(x, leak * x)


def batch_moments(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``tf.nn.moments(x, [0,1,2])``: per-channel mean and *biased* variance (distriubted_model.py:37)."""
    dims = tuple(range(x.dim() - 1))
    mean = x.mean(dim=dims)
    var = ((x - mean) ** 2).mean(dim=dims)
    return mean, var


def batch_norm_apply(x: torch.Tensor, mean: torch.Tensor, var: torch.Tensor, beta: torch.Tensor,
                     gamma: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    """``batch_norm_with_global_normalization(.., scale_after_normalization=True)``
    (distriubted_model.py:49-50): ``gamma * (x - mean) * rsqrt(var + eps) + beta``."""
    return (x - mean) * (gamma * torch.rsqrt(var + eps)) + beta


def ema_update_(shadow: torch.Tensor, value: torch.Tensor, decay: float = 0.9) -> None:
    """``shadow -= (1 - decay) * (shadow - value)`` (tf.train.ExponentialMovingAverage.apply,
    distriubted_model.py:41)."""
    shadow.sub_((1.0 - decay) * (shadow - value.detach().to(shadow.dtype)))


def bce_with_logits_mean(logits: torch.Tensor, target: float) -> torch.Tensor:
    """``reduce_mean(sigmoid_cross_entropy_with_logits(logits, target))`` (image_train.py:91-95):
    ``max(x,0) - x*t + log1p(exp(-|x|))``."""
    x = logits.float()
    return (torch.clamp(x, min=0) - x * target + torch.log1p(torch.exp(-x.abs()))).mean()


def zero_fraction(x: torch.Tensor) -> torch.Tensor:
    """``tf.nn.zero_fraction`` -- the per-layer "sparsity" summary (distriubted_model.py:80)."""
    return (x == 0).float().mean()
