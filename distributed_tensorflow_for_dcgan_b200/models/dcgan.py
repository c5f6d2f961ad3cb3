"""DCGAN generator / discriminator / sampler as pure functions over a TF-named parameter dict.

Architecture and call order follow the reference ``generator`` / ``discriminator`` / ``sampler``
(``distriubted_model.py:83-153``).  This module is the oracle and the CPU path; the B200 path
(``engine.py``) runs the same math on hand-written kernels from the very same parameter tensors.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Tuple

import torch

from ..config import ModelConfig
from ..ops import functional as Fn
from . import manifest

Params = Dict[str, torch.Tensor]
# hook(scope, batch_mean, batch_var): receives batch moments of every training-mode BN call
BNHook = Optional[Callable[[str, torch.Tensor, torch.Tensor], None]]
ActHook = Optional[Callable[[str, torch.Tensor], None]]


def _bn_train(x, p: Params, name: str, scope: str, cfg: ModelConfig, hook: BNHook):
    mean, var = Fn.batch_moments(x)
    if hook is not None:
        hook(scope, mean, var)
    return Fn.batch_norm_apply(x, mean, var, p[f"{name}/beta"], p[f"{name}/gamma"], cfg.bn_eps)


def _bn_infer(x, p: Params, name: str, ema: Params, scope: str, cfg: ModelConfig):
    m, v = manifest.ema_names(scope)
    return Fn.batch_norm_apply(x, ema[m], ema[v], p[f"{name}/beta"], p[f"{name}/gamma"], cfg.bn_eps)


def generator(z: torch.Tensor, p: Params, cfg: ModelConfig, bn_hook: BNHook = None,
              act_hook: ActHook = None) -> torch.Tensor:
    """``generator(z)`` (distriubted_model.py:83-111): linear -> [N,s16,s16,8gf] -> (BN, ReLU,
    deconv) x4 -> tanh.  Returns NHWC images in [-1, 1]."""
    s16 = cfg.s16
    h = Fn.linear(z, p["g_h0_lin/Matrix"], p["g_h0_lin/bias"]).reshape(-1, s16, s16, cfg.gf_dim * 8)
    h = torch.relu(_bn_train(h, p, "g_bn0", "g_bn0", cfg, bn_hook))
    if act_hook:
        act_hook("g_h0_relu", h)
    for i in (1, 2, 3):
        h = Fn.deconv2d(h, p[f"g_h{i}/w"], p[f"g_h{i}/biases"])
        h = torch.relu(_bn_train(h, p, f"g_bn{i}", f"g_bn{i}", cfg, bn_hook))
        if act_hook:
            act_hook(f"g_h{i}_relu", h)
    h = Fn.deconv2d(h, p["g_h4/w"], p["g_h4/biases"])
    if act_hook:
        act_hook("g_h4", h)
    return torch.tanh(h)


def sampler(z: torch.Tensor, p: Params, ema: Params, cfg: ModelConfig) -> torch.Tensor:
    """``sampler(z)`` (distriubted_model.py:131-153): the generator with ``train=False`` batch
    norm, i.e. normalising with the EMA shadows of the generator's training-mode moments."""
    s16 = cfg.s16
    h = Fn.linear(z, p["g_h0_lin/Matrix"], p["g_h0_lin/bias"]).reshape(-1, s16, s16, cfg.gf_dim * 8)
    h = torch.relu(_bn_infer(h, p, "g_bn0", ema, "g_bn0", cfg))
    for i in (1, 2, 3):
        h = Fn.deconv2d(h, p[f"g_h{i}/w"], p[f"g_h{i}/biases"])
        h = torch.relu(_bn_infer(h, p, f"g_bn{i}", ema, f"g_bn{i}", cfg))
    return torch.tanh(Fn.deconv2d(h, p["g_h4/w"], p["g_h4/biases"]))


def discriminator(image: torch.Tensor, p: Params, cfg: ModelConfig, second_call: bool = False,
                  bn_hook: BNHook = None, act_hook: ActHook = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """``discriminator(image, reuse)`` (distriubted_model.py:114-128).  Returns
    ``(sigmoid(logits), logits)``.  ``second_call`` selects the EMA scope of the fake-batch call
    (the reference builds D twice: on real images, then with ``reuse=True`` on G's output)."""
    sfx = "_1" if second_call else ""
    h = Fn.lrelu(Fn.conv2d_same(image, p["d_h0_conv/w"], p["d_h0_conv/biases"]), cfg.lrelu_leak)
    if act_hook:
        act_hook("d_h0_conv" + sfx, h)
    for i in (1, 2, 3):
        h = Fn.conv2d_same(h, p[f"d_h{i}_conv/w"], p[f"d_h{i}_conv/biases"])
        h = Fn.lrelu(_bn_train(h, p, f"d_bn{i}", f"d_bn{i}{sfx}", cfg, bn_hook), cfg.lrelu_leak)
        if act_hook:
            act_hook(f"d_h{i}_conv" + sfx, h)
    logits = Fn.linear(h.reshape(h.shape[0], -1), p["d_h3_lin/Matrix"], p["d_h3_lin/bias"])
    if act_hook:
        act_hook("d_h3_lin" + sfx, logits)
    return torch.sigmoid(logits), logits


def gan_losses(d_logits_real: torch.Tensor, d_logits_fake: torch.Tensor):
    """(d_loss_real, d_loss_fake, g_loss) -- image_train.py:91-96."""
    return (Fn.bce_with_logits_mean(d_logits_real, 1.0),
            Fn.bce_with_logits_mean(d_logits_fake, 0.0),
            Fn.bce_with_logits_mean(d_logits_fake, 1.0))


def init_params(cfg: ModelConfig, seed: int = 0, device="cpu", dtype=torch.float32) -> Params:
    """Initialisers of SURVEY.md 2.5: N(0,0.02) / truncated N(0,0.02) (two-sigma re-draw, as
    ``tf.truncated_normal_initializer``) / zeros / N(1,0.02) for gamma."""
    gen = torch.Generator(device="cpu").manual_seed(seed)
    out: Params = {}
    for spec in manifest.trainable_specs(cfg):
        if spec.init == "zeros":
            t = torch.zeros(spec.shape)
        elif spec.init == "normal":
            t = torch.randn(spec.shape, generator=gen) * spec.stddev
        elif spec.init == "normal_one":
            t = 1.0 + torch.randn(spec.shape, generator=gen) * spec.stddev
        elif spec.init == "trunc_normal":
            t = torch.empty(spec.shape)
            torch.nn.init.trunc_normal_(t, 0.0, spec.stddev, -2 * spec.stddev, 2 * spec.stddev, generator=gen)
        else:
            raise ValueError(spec.init)
        out[spec.name] = t.to(device=device, dtype=dtype)
    return out


def init_ema(cfg: ModelConfig, device="cpu") -> Params:
    """EMA shadows of tensors start at zero (tf.train.ExponentialMovingAverage on non-Variables)."""
    return {s.name: torch.zeros(s.shape, device=device) for s in manifest.ema_specs(cfg)}
