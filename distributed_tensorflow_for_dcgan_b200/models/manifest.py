"""Variable manifest = the checkpoint name contract.

Names follow the reference's ``tf.variable_scope(name)`` + ``get_variable`` calls
(``distriubted_model.py:30-34,163-168,179-182,193-197``).  Layouts are TensorFlow's:
conv ``[kh,kw,Cin,Cout]`` (HWIO), deconv ``[kh,kw,Cout,Cin]``, linear ``[in,out]``.
Creation order is generator first, then discriminator (``image_train.py:79-82``).
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Tuple

from ..config import ModelConfig


@dataclasses.dataclass(frozen=True)
class VarSpec:
    name: str
    shape: Tuple[int, ...]
    init: str          # "normal", "trunc_normal", "zeros", "normal_one" (N(1, 0.02))
    stddev: float = 0.02

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


def generator_specs(cfg: ModelConfig) -> List[VarSpec]:
    gf, c = cfg.gf_dim, cfg.c_dim
    specs = [
        VarSpec("g_h0_lin/Matrix", (cfg.z_dim, cfg.g_h0_features), "normal"),
        VarSpec("g_h0_lin/bias", (cfg.g_h0_features,), "zeros"),
        VarSpec("g_bn0/beta", (gf * 8,), "zeros"),
        VarSpec("g_bn0/gamma", (gf * 8,), "normal_one"),
    ]
    chans = [gf * 8, gf * 4, gf * 2, gf, c]
    for i in range(1, 5):
        cin, cout = chans[i - 1], chans[i]
        # deconv filter is [kh, kw, output_channels, in_channels] (distriubted_model.py:193-195)
        specs.append(VarSpec(f"g_h{i}/w", (5, 5, cout, cin), "normal"))
        specs.append(VarSpec(f"g_h{i}/biases", (cout,), "zeros"))
        if i < 4:
            specs.append(VarSpec(f"g_bn{i}/beta", (cout,), "zeros"))
            specs.append(VarSpec(f"g_bn{i}/gamma", (cout,), "normal_one"))
    return specs


def discriminator_specs(cfg: ModelConfig) -> List[VarSpec]:
    df, c = cfg.df_dim, cfg.c_dim
    chans = [c, df, df * 2, df * 4, df * 8]
    specs: List[VarSpec] = []
    for i in range(4):
        cin, cout = chans[i], chans[i + 1]
        specs.append(VarSpec(f"d_h{i}_conv/w", (5, 5, cin, cout), "trunc_normal"))
        specs.append(VarSpec(f"d_h{i}_conv/biases", (cout,), "zeros"))
        if i > 0:  # d_bn0 exists in the reference but is never applied (distriubted_model.py:55,118)
            specs.append(VarSpec(f"d_bn{i}/beta", (cout,), "zeros"))
            specs.append(VarSpec(f"d_bn{i}/gamma", (cout,), "normal_one"))
    specs.append(VarSpec("d_h3_lin/Matrix", (cfg.d_flat_features, 1), "normal"))
    specs.append(VarSpec("d_h3_lin/bias", (1,), "zeros"))
    return specs


def trainable_specs(cfg: ModelConfig) -> List[VarSpec]:
    return generator_specs(cfg) + discriminator_specs(cfg)


def is_d_var(name: str) -> bool:
    """``'d_' in var.name`` (image_train.py:107)."""
    return "d_" in name


def is_g_var(name: str) -> bool:
    """``'g_' in var.name`` (image_train.py:108)."""
    return "g_" in name


def bn_sites(cfg: ModelConfig) -> List[Tuple[str, int]]:
    """(EMA scope, channels) for every training-mode batch_norm call, in graph order.

    The reference calls g_bn0..3 once and d_bn1..3 twice (real batch, then fake batch), and each
    call owns its own pair of EMA shadows.  TensorFlow derives the shadow names
    from internal op names, which cannot be recovered from the tree; the canonical names chosen
    here are ``<scope>/moments/{mean,variance}/ExponentialMovingAverage`` where the second call
    of a scope is ``<bn>_1`` (the name scope TensorFlow opens on re-entry).
    """
    gf, df = cfg.gf_dim, cfg.df_dim
    sites = [("g_bn0", gf * 8), ("g_bn1", gf * 4), ("g_bn2", gf * 2), ("g_bn3", gf)]
    for suffix in ("", "_1"):
        sites += [(f"d_bn1{suffix}", df * 2), (f"d_bn2{suffix}", df * 4), (f"d_bn3{suffix}", df * 8)]
    return sites


def ema_names(scope: str) -> Tuple[str, str]:
    return (f"{scope}/moments/mean/ExponentialMovingAverage",
            f"{scope}/moments/variance/ExponentialMovingAverage")


def ema_specs(cfg: ModelConfig) -> List[VarSpec]:
    out = []
    for scope, ch in bn_sites(cfg):
        m, v = ema_names(scope)
        out.append(VarSpec(m, (ch,), "zeros"))
        out.append(VarSpec(v, (ch,), "zeros"))
    return out


GLOBAL_STEP_NAME = "Variable"  # the unnamed tf.Variable(0) at image_train.py:68


def adam_slot_names(var_name: str) -> Tuple[str, str]:
    """TensorFlow slot naming.  The reference's Saver does not contain them (it is built before
    ``minimize``, image_train.py:103 vs :109-112 -- a latent bug we do not replicate); we store them
    under these extra keys so that the weight subset still maps one-to-one."""
    return f"{var_name}/Adam", f"{var_name}/Adam_1"


def param_counts(cfg: ModelConfig) -> Dict[str, int]:
    g = sum(s.numel for s in generator_specs(cfg))
    d = sum(s.numel for s in discriminator_specs(cfg))
    return {"G": g, "D": d, "total": g + d}
