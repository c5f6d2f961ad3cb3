"""Tensor-level wrappers over the sm_100a kernels (one function per C entry point).

All functions launch on the current CUDA stream and allocate nothing unless an output tensor is
omitted.  Activations are NHWC bf16; see include/dcgan_b200.h for layouts.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib, autotune
from ._lib import call, ptr

ACT = {"none": 0, "relu": 1, "lrelu": 2, "tanh": 3, "sigmoid": 4}


def _bf16(shape, like):
    return torch.empty(shape, dtype=torch.bfloat16, device=like.device)


def conv_fprop(x, w_nk, bias=None, act="none", out=None, block_n=0):
    """5x5 s2 SAME conv.  x [N,H,W,Cin] bf16, w_nk [25,Cout,Cin] bf16 -> [N,H/2,W/2,Cout] bf16."""
    N, H, W, Cin = x.shape
    Cout = w_nk.shape[1]
    if block_n == 0:
        autotune.ensure(autotune.KIND_F, N, H, W, Cin, Cout)
    if out is None:
        out = _bf16((N, H // 2, W // 2, Cout), x)
    call("dcgan_conv_fprop", ptr(x), ptr(w_nk), ptr(bias), ptr(out), N, H, W, Cin, Cout, ACT[act], block_n)
    return out


def conv_fprop_kn(x, w_kn, bias=None, act="none", out=None, block_n=0):
    """5x5 s2 SAME conv with the filter in TF layout: w_kn [25,Cin,Cout] bf16 (no transposed copy)."""
    N, H, W, Cin = x.shape
    Cout = w_kn.shape[2]
    if block_n == 0:
        autotune.ensure(autotune.KIND_F, N, H, W, Cin, Cout)
    if out is None:
        out = _bf16((N, H // 2, W // 2, Cout), x)
    call("dcgan_conv_fprop_kn", ptr(x), ptr(w_kn), ptr(bias), ptr(out), N, H, W, Cin, Cout, ACT[act], block_n)
    return out


def conv_tconv(x, w_nk, bias=None, act="none", out=None, block_n=0):
    """5x5 s2 transposed conv (TF crop).  x [N,H,W,Cin], w_nk [25,Cout,Cin] -> [N,2H,2W,Cout]."""
    N, H, W, Cin = x.shape
    Cout = w_nk.shape[1]
    if block_n == 0:
        autotune.ensure(autotune.KIND_T, N, H, W, Cin, Cout)
    if out is None:
        out = _bf16((N, 2 * H, 2 * W, Cout), x)
    call("dcgan_conv_tconv", ptr(x), ptr(w_nk), ptr(bias), ptr(out), N, H, W, Cin, Cout, ACT[act], block_n)
    return out


def conv_fprop_kn_stats(x, w_kn, bias, out, stats, groups, block_n=0):
    """conv_fprop_kn + batch-norm moments of the output accumulated into stats [groups,2,Cout]."""
    N, H, W, Cin = x.shape
    if block_n == 0:
        autotune.ensure(autotune.KIND_F, N, H, W, Cin, w_kn.shape[2])
    call("dcgan_conv_fprop_kn_stats", ptr(x), ptr(w_kn), ptr(bias), ptr(out), ptr(stats), groups, N, H, W, Cin,
         w_kn.shape[2], block_n)
    return out


def conv_tconv_stats(x, w_nk, bias, out, stats, groups, block_n=0):
    N, H, W, Cin = x.shape
    if block_n == 0:
        autotune.ensure(autotune.KIND_T, N, H, W, Cin, w_nk.shape[1])
    call("dcgan_conv_tconv_stats", ptr(x), ptr(w_nk), ptr(bias), ptr(out), ptr(stats), groups, N, H, W, Cin,
         w_nk.shape[1], block_n)
    return out


def conv_dgrad_bnred(x, w, out, pre, saved, gamma, beta, red, slabs, stat_groups, act, transposed, block_n=0):
    """Data-gradient conv (transposed: w [25,Cout,Cin]; strided: w [25,Cin,Cout]) whose epilogue also accumulates
    the batch-norm backward reductions ``red [slabs,2,Cout]`` of the layer that ``out`` is the activation gradient
    of (``pre`` / ``saved`` / ``gamma`` / ``beta`` are that layer's)."""
    N, H, W, Cin = x.shape
    if block_n == 0:
        autotune.ensure(autotune.KIND_T if transposed else autotune.KIND_F, N, H, W, Cin, w.shape[1] if transposed else w.shape[2])
    Cout = w.shape[1] if transposed else w.shape[2]
    call("dcgan_conv_tconv_bnred" if transposed else "dcgan_conv_fprop_kn_bnred", ptr(x), ptr(w), ptr(out), ptr(pre),
         ptr(saved), ptr(gamma), ptr(beta), ptr(red), slabs, stat_groups, ACT[act], N, H, W, Cin, Cout, block_n)
    return out


def conv_dgrad_bnbwd(x, w, out, pre, saved, gamma, beta, red, dpre, dgamma, dbeta, slabs, stat_groups, param_slabs,
                     act, transposed, block_n=0):
    """``conv_dgrad_bnred`` with the rest of that layer's batch-norm backward pass in the same launch: behind a
    grid-wide barrier the kernel turns ``out`` (= da) into ``dpre`` and writes ``dgamma`` / ``dbeta``.  Kernels of
    this family (this, ``conv_bn_fwd``, ``bn_forward_fused``, ``bn_backward_fused``) go on ONE stream per device."""
    N, H, W, Cin = x.shape
    if block_n == 0:
        autotune.ensure(autotune.KIND_T if transposed else autotune.KIND_F, N, H, W, Cin, w.shape[1] if transposed else w.shape[2])
    Cout = w.shape[1] if transposed else w.shape[2]
    call("dcgan_conv_tconv_bnbwd" if transposed else "dcgan_conv_fprop_kn_bnbwd", ptr(x), ptr(w), ptr(out), ptr(pre),
         ptr(saved), ptr(gamma), ptr(beta), ptr(red), ptr(dpre), ptr(dgamma), ptr(dbeta), slabs, stat_groups,
         param_slabs, ACT[act], N, H, W, Cin, Cout, block_n)
    return dpre


def conv_tconv_actbwd(x, w_nk, out, act_saved, dbias, slabs, stat_groups, bias_slabs, act, block_n=0):
    """Transposed conv (data gradient of a strided conv) whose epilogue applies the activation backward of the layer
    below (bias + activation, no batch norm) and accumulates that layer's bias gradient: ``conv_tconv`` + ``act_bwd``."""
    N, H, W, Cin = x.shape
    Cout = w_nk.shape[1]
    if block_n == 0:
        autotune.ensure(autotune.KIND_T, N, H, W, Cin, Cout)
    call("dcgan_conv_tconv_actbwd", ptr(x), ptr(w_nk), ptr(out), ptr(act_saved), ptr(dbias), slabs, stat_groups,
         bias_slabs, ACT[act], N, H, W, Cin, Cout, block_n)
    return out


def conv_bn_fwd(x, w, bias, pre, act_out, sums, gamma, beta, saved, ema_mean, ema_var, groups, act, transposed,
                eps=1e-5, decay=0.9, block_n=0):
    """Forward conv (transposed: w [25,Cout,Cin]; strided: w [25,Cin,Cout]) + training-mode batch norm + activation
    in one launch: ``pre`` (bf16 conv output, kept for backward), ``sums`` [groups,2,Cout] (+=), then
    ``act_out = act(bn(pre))``, ``saved`` = (mean, rstd) and the EMA shadows."""
    N, H, W, Cin = x.shape
    if block_n == 0:
        autotune.ensure(autotune.KIND_T if transposed else autotune.KIND_F, N, H, W, Cin, w.shape[1] if transposed else w.shape[2])
    Cout = w.shape[1] if transposed else w.shape[2]
    call("dcgan_conv_tconv_bn" if transposed else "dcgan_conv_fprop_kn_bn", ptr(x), ptr(w), ptr(bias), ptr(pre),
         ptr(act_out), ptr(sums), ptr(gamma), ptr(beta), ptr(saved), ptr(ema_mean), ptr(ema_var), groups, N, H, W,
         Cin, Cout, eps, decay, ACT[act], block_n)
    return act_out


def conv_wgrad(big, small, dw, splits=0, block_n=0):
    """dw[25,Cg,Cd] (f32, +=) from big [N,2Hs,2Ws,Cg] and small [N,Hs,Ws,Cd]."""
    N, Hs, Ws, Cd = small.shape
    Cg = big.shape[3]
    if block_n == 0 and splits == 0:
        autotune.ensure(autotune.KIND_W, N, Hs, Ws, Cg, Cd)
    call("dcgan_conv_wgrad", ptr(big), ptr(small), ptr(dw), N, Hs, Ws, Cg, Cd, splits, block_n)
    return dw


def _tc_width_ok(w_out: int, h_out: int) -> bool:
    return w_out in (32, 64, 128) and (h_out * w_out) % 128 == 0


def c3_fprop(x, w, bias=None, act="none", out=None, tensor_cores=True):
    """3 -> C strided conv; tensor-core kernel when the output width allows, CUDA-core kernel otherwise."""
    N, H, W, _ = x.shape
    C = w.shape[-1]
    if out is None:
        out = _bf16((N, H // 2, W // 2, C), x)
    if tensor_cores and _tc_width_ok(W // 2, H // 2):
        call("dcgan_c3_fprop_tc", ptr(x), ptr(w), ptr(bias), ptr(out), N, H, W, C, ACT[act])
    else:
        call("dcgan_c3_fprop", ptr(x), ptr(w), ptr(bias), ptr(out), N, H, W, C, ACT[act])
    return out


def c3_tconv(x, w, bias=None, act="none", aux=None, out=None, colsum=None, w_bf16=None):
    """mode 0: act(v + bias); with ``aux``: v * (1 - aux^2).  ``colsum`` (f32[3]) += channel sums.
    With ``w_bf16`` (the bf16 shadow of the [25,3,C] filter) and C in (64, 128) the tensor-core kernel runs."""
    N, H, W, C = x.shape
    if out is None:
        out = _bf16((N, 2 * H, 2 * W, 3), x)
    if w_bf16 is not None and C in (64, 128) and H % 8 == 0 and W % 16 == 0:
        call("dcgan_c3_tconv_tc", ptr(x), ptr(w_bf16), ptr(bias), ptr(aux), ptr(out), ptr(colsum), N, H, W, C,
             ACT[act], 0 if aux is None else 1)
    else:
        call("dcgan_c3_tconv", ptr(x), ptr(w), ptr(bias), ptr(aux), ptr(out), ptr(colsum), N, H, W, C, ACT[act],
             0 if aux is None else 1)
    return out


def c3_wgrad(x3, y, dw, tensor_cores=True):
    N, Hs, Ws, C = y.shape
    if tensor_cores and _tc_width_ok(Ws, Hs):
        call("dcgan_c3_wgrad_tc", ptr(x3), ptr(y), ptr(dw), N, Hs, Ws, C)
        _lib.launch_count += 1  # two kernels: per-CTA partial tiles, then the ordered reduce
    else:
        call("dcgan_c3_wgrad", ptr(x3), ptr(y), ptr(dw), N, Hs, Ws, C)
    return dw


def bn_stats(x2d, sums, groups):
    rows, C = x2d.shape[0] // groups, x2d.shape[-1]
    call("dcgan_bn_stats", ptr(x2d), ptr(sums), groups, rows, C)
    return sums


def bn_apply(x2d, y2d, sums, gamma, beta, saved, ema_mean, ema_var, groups, act, eps=1e-5, decay=0.9,
             use_ema=False):
    rows, C = x2d.shape[0] // groups, x2d.shape[-1]
    call("dcgan_bn_apply", ptr(x2d), ptr(y2d), ptr(sums), ptr(gamma), ptr(beta), ptr(saved), ptr(ema_mean),
         ptr(ema_var), groups, rows, C, eps, decay, ACT[act], int(use_ema))
    return y2d


def bn_bwd_reduce(da2d, pre2d, saved, gamma, beta, red, groups, stat_groups, act):
    rows, C = da2d.shape[0] // groups, da2d.shape[-1]
    call("dcgan_bn_bwd_reduce", ptr(da2d), ptr(pre2d), ptr(saved), ptr(gamma), ptr(beta), ptr(red), groups,
         stat_groups, rows, C, ACT[act])
    return red


def bn_bwd_apply(da2d, pre2d, dpre2d, saved, gamma, beta, red, dgamma, dbeta, groups, stat_groups,
                 param_groups, act):
    rows, C = da2d.shape[0] // groups, da2d.shape[-1]
    call("dcgan_bn_bwd_apply", ptr(da2d), ptr(pre2d), ptr(dpre2d), ptr(saved), ptr(gamma), ptr(beta), ptr(red),
         ptr(dgamma), ptr(dbeta), groups, stat_groups, param_groups, rows, C, ACT[act])
    return dpre2d


def bn_forward_fused(x2d, y2d, sums, gamma, beta, saved, ema_mean, ema_var, barrier, groups, act, eps=1e-5,
                     decay=0.9):
    """``bn_stats`` + ``bn_apply`` (training mode) in one launch; ``barrier`` = two zero int32 on the device."""
    rows, C = x2d.shape[0] // groups, x2d.shape[-1]
    call("dcgan_bn_forward_fused", ptr(x2d), ptr(y2d), ptr(sums), ptr(gamma), ptr(beta), ptr(saved), ptr(ema_mean),
         ptr(ema_var), ptr(barrier), groups, rows, C, eps, decay, ACT[act])
    return y2d


def bn_backward_fused(da2d, pre2d, dpre2d, saved, gamma, beta, red, dgamma, dbeta, barrier, groups, stat_groups,
                      param_groups, act):
    """``bn_bwd_reduce`` + ``bn_bwd_apply`` in one launch."""
    rows, C = da2d.shape[0] // groups, da2d.shape[-1]
    call("dcgan_bn_backward_fused", ptr(da2d), ptr(pre2d), ptr(dpre2d), ptr(saved), ptr(gamma), ptr(beta), ptr(red),
         ptr(dgamma), ptr(dbeta), ptr(barrier), groups, stat_groups, param_groups, rows, C, ACT[act])
    return dpre2d


def act_bwd(da2d, a2d, dpre2d, dbias, act, bias_rows):
    rows, C = da2d.shape
    call("dcgan_act_bwd", ptr(da2d), ptr(a2d), ptr(dpre2d), ptr(dbias), rows, a2d.shape[0], C, ACT[act], bias_rows)
    return dpre2d


def lin_fwd(z, w_bf16, bias, out):
    NB, K = z.shape
    F = w_bf16.shape[1]
    call("dcgan_lin_fwd", ptr(z), ptr(w_bf16), ptr(bias), ptr(out), NB, K, F)
    return out


def lin_wgrad(z, dy, dW, dbias=None):
    NB, K = z.shape
    F = dW.shape[1]
    call("dcgan_lin_wgrad", ptr(z), ptr(dy), ptr(dW), ptr(dbias), NB, K, F)
    return dW


def head_fwd(h2d, w, b, logits):
    rows, F = h2d.shape
    call("dcgan_head_fwd", ptr(h2d), ptr(w), ptr(b), ptr(logits), rows, F)
    return logits


def gan_loss(logits, losses, dlogit, probs, counters, B, grad_div=1.0, advance=True):
    call("dcgan_gan_loss", ptr(logits), ptr(losses), ptr(dlogit), ptr(probs), ptr(counters), B, float(grad_div),
         int(advance))


def head_bwd(h2d, w, dlogit, dh, dW, db, B):
    F = h2d.shape[1]
    call("dcgan_head_bwd", ptr(h2d), ptr(w), ptr(dlogit), ptr(dh), ptr(dW), ptr(db), B, F)


def head_loss_bwd(h2d, w, logits, losses, dlogit, probs, counters, dh, dW, db, B, grad_div=1.0, advance=True):
    """gan_loss + head_bwd in one launch."""
    F = h2d.shape[1]
    call("dcgan_head_loss_bwd", ptr(h2d), ptr(w), ptr(logits), ptr(losses), ptr(dlogit), ptr(probs), ptr(counters),
         ptr(dh), ptr(dW), ptr(db), B, F, float(grad_div), int(advance))


def adam(p, g, m, v, p_bf16, counters, lr=2e-4, b1=0.5, b2=0.999, eps=1e-8, gscale=1.0):
    call("dcgan_adam", ptr(p), ptr(g), ptr(m), ptr(v), ptr(p_bf16), ptr(counters), p.numel(), lr, b1, b2, eps,
         gscale)


def cast_bf16(src, dst):
    call("dcgan_cast_bf16", ptr(src), ptr(dst), src.numel())
    return dst


def transpose_taps(src, dst, taps, R, C):
    call("dcgan_transpose_taps", ptr(src), ptr(dst), taps, R, C)
    return dst


def colsum(x2d, out):
    rows, C = x2d.shape
    call("dcgan_colsum", ptr(x2d), ptr(out), rows, C)
    return out


def zero_(t):
    """Stream-ordered zero fill of a contiguous tensor (a memset node under graph capture, not a fill kernel)."""
    assert t.is_contiguous()
    call("dcgan_zero_async", ptr(t), t.numel() * t.element_size())
    _lib.launch_count -= 1  # not a kernel launch
    return t


def decode_f64(src, dst, scale=1.0, shift=0.0):
    """float64 records -> bf16 (``dst`` bf16) or float32 (``dst`` float32), ``x * scale + shift`` on the way."""
    call("dcgan_decode_f64_f32" if dst.dtype == torch.float32 else "dcgan_decode_f64", ptr(src), ptr(dst), src.numel(),
         scale, shift)
    return dst


def f32_to_bf16(src, dst, scale=1.0, shift=0.0):
    call("dcgan_f32_to_bf16", ptr(src), ptr(dst), src.numel(), scale, shift)
    return dst


def grid_merge(img, cols=8):
    N, H, W, C = img.shape
    rows = (N + cols - 1) // cols
    out = torch.zeros((rows * H, cols * W, C), dtype=torch.float32, device=img.device)
    call("dcgan_grid_merge", ptr(img), ptr(out), N, H, W, C, cols)
    return out


def zero_fraction(x):
    cnt = torch.zeros(1, dtype=torch.int64, device=x.device)
    call("dcgan_count_zeros", ptr(x), ptr(cnt), x.numel())
    return cnt.float() / x.numel()


def check_device_error():
    e = _lib.last_device_error()
    if e:
        raise _lib.KernelError(f"device-side error flag = {e} (1: mbarrier wait timed out, 2: peer wait timed out)")
