"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op
(tier "Unit / op parity")."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from distributed_tensorflow_for_dcgan_b200.ops import functional as Fn


def _rand(shape, scale=1.0, dtype=torch.bfloat16):
    return (torch.randn(shape, device="cuda") * scale).to(dtype)


def _close(got, ref, rel=1e-2, name=""):
    got, ref = got.float(), ref.float()
    err = (got - ref).abs().max().item()
    den = ref.abs().max().item()
    assert den > 0, name
    assert err <= rel * den, f"{name}: max err {err} vs max ref {den}"


def _w_hwio_to_nk(w_hwio):  # [5,5,Cin,Cout] -> [25,Cout,Cin]
    return w_hwio.permute(0, 1, 3, 2).reshape(25, w_hwio.shape[3], w_hwio.shape[2]).contiguous()


@pytest.mark.parametrize("N,H,Cin,Cout", [(2, 16, 64, 128), (4, 8, 128, 64), (1, 8, 64, 256)])
def test_conv_fprop(cuda_lib, N, H, Cin, Cout):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    x = _rand((N, H, H, Cin))
    w = _rand((5, 5, Cin, Cout), 0.05)
    b = torch.randn(Cout, device="cuda") * 0.1
    out = K.conv_fprop(x, _w_hwio_to_nk(w), b, act="lrelu")
    ref = Fn.lrelu(Fn.conv2d_same(x.float(), w.float(), b))
    K.check_device_error()
    _close(out, ref, name="conv_fprop")


@pytest.mark.parametrize("N,H,Cin,Cout", [(2, 16, 64, 128), (4, 8, 128, 64), (1, 8, 64, 256), (2, 8, 256, 512)])
def test_conv_fprop_tf_layout(cuda_lib, N, H, Cin, Cout):
    """Filter consumed in its HWIO layout through an MN-major B operand (no transposed copy)."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    x = _rand((N, H, H, Cin))
    w = _rand((5, 5, Cin, Cout), 0.05)
    b = torch.randn(Cout, device="cuda") * 0.1
    out = K.conv_fprop_kn(x, w.view(25, Cin, Cout), b, act="lrelu")
    ref = Fn.lrelu(Fn.conv2d_same(x.float(), w.float(), b))
    K.check_device_error()
    _close(out, ref, name="conv_fprop_kn")


def test_conv_epilogue_batchnorm_moments(cuda_lib):
    """The conv epilogue's fused (sum, sum-of-squares) equal the moments of the conv output, per group."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    N, H, Cin, Cout, G = 8, 16, 64, 128, 2
    x = _rand((N, H, H, Cin))
    w = _rand((5, 5, Cin, Cout), 0.05)
    b = torch.randn(Cout, device="cuda") * 0.1
    out = torch.empty(N, H // 2, H // 2, Cout, dtype=torch.bfloat16, device="cuda")
    stats = torch.zeros(G, 2, Cout, device="cuda")
    K.conv_fprop_kn_stats(x, w.view(25, Cin, Cout), b, out, stats, G)
    ref = Fn.conv2d_same(x.float(), w.float(), b).view(G, -1, Cout)
    _close(out.view(G, -1, Cout), ref, name="fprop_stats out")
    _close(stats[:, 0], ref.sum(1), rel=5e-3, name="fused sum")
    _close(stats[:, 1], (ref * ref).sum(1), rel=5e-3, name="fused sumsq")
    # transposed mode, one group
    xt = _rand((4, 8, 8, 128))
    wt = _rand((5, 5, 64, 128), 0.05)
    outt = torch.empty(4, 16, 16, 64, dtype=torch.bfloat16, device="cuda")
    st = torch.zeros(1, 2, 64, device="cuda")
    K.conv_tconv_stats(xt, wt.view(25, 64, 128), None, outt, st, 1)
    reft = Fn.deconv2d(xt.float(), wt.float()).reshape(-1, 64)
    _close(outt.view(-1, 64), reft, name="tconv_stats out")
    _close(st[0, 0], reft.sum(0), rel=5e-3, name="tconv fused sum")
    _close(st[0, 1], (reft * reft).sum(0), rel=5e-3, name="tconv fused sumsq")
    K.check_device_error()


@pytest.mark.parametrize("N,H,Cin,Cout", [(2, 8, 64, 64), (2, 4, 128, 128), (1, 8, 256, 128)])
def test_conv_tconv(cuda_lib, N, H, Cin, Cout):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    x = _rand((N, H, H, Cin))
    w = _rand((5, 5, Cout, Cin), 0.05)  # TF deconv layout [kh,kw,Cout,Cin]
    b = torch.randn(Cout, device="cuda") * 0.1
    out = K.conv_tconv(x, w.reshape(25, Cout, Cin), b)
    ref = Fn.deconv2d(x.float(), w.float(), b)
    K.check_device_error()
    _close(out, ref, name="conv_tconv")


@pytest.mark.parametrize("N,Hs,Cg,Cd", [(2, 8, 64, 128), (4, 4, 256, 256), (4, 8, 128, 256)])
def test_conv_wgrad_conv_layer(cuda_lib, N, Hs, Cg, Cd):
    """conv2d filter gradient: big = layer input, small = output gradient -> HWIO."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    x = _rand((N, 2 * Hs, 2 * Hs, Cg))
    dy = _rand((N, Hs, Hs, Cd))
    dw = torch.zeros(25, Cg, Cd, device="cuda")
    K.conv_wgrad(x, dy, dw)
    w = torch.zeros(5, 5, Cg, Cd, device="cuda", requires_grad=True)
    y = Fn.conv2d_same(x.float(), w)
    ref, = torch.autograd.grad(y, w, dy.float())
    K.check_device_error()
    _close(dw.view(5, 5, Cg, Cd), ref, rel=2e-3, name="conv_wgrad")


def test_conv_wgrad_deconv_layer(cuda_lib):
    """conv2d_transpose filter gradient: big = output gradient, small = layer input -> [kh,kw,Cout,Cin]."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    N, Hs, Cout, Cin = 2, 8, 64, 128
    x = _rand((N, Hs, Hs, Cin))
    dy = _rand((N, 2 * Hs, 2 * Hs, Cout))
    dw = torch.zeros(25, Cout, Cin, device="cuda")
    K.conv_wgrad(dy, x, dw)
    w = torch.zeros(5, 5, Cout, Cin, device="cuda", requires_grad=True)
    y = Fn.deconv2d(x.float(), w)
    ref, = torch.autograd.grad(y, w, dy.float())
    K.check_device_error()
    _close(dw.view(5, 5, Cout, Cin), ref, rel=2e-3, name="deconv_wgrad")


@pytest.mark.parametrize("tc", [True, False])
@pytest.mark.parametrize("S,C", [(64, 64), (128, 128)])
def test_c3_fprop(cuda_lib, tc, S, C):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    x = _rand((3, S, S, 3))
    w = torch.randn(5, 5, 3, C, device="cuda") * 0.05
    b = torch.randn(C, device="cuda") * 0.1
    out = K.c3_fprop(x, w, b, act="lrelu", tensor_cores=tc)
    ref = Fn.lrelu(Fn.conv2d_same(x.float(), w.to(torch.bfloat16).float() if tc else w, b))
    K.check_device_error()
    _close(out, ref, name="c3_fprop")


@pytest.mark.parametrize("C", [64, 128])  # 128 = the wide configuration: also on the tensor cores
@pytest.mark.parametrize("tc", [True, False])
def test_c3_tconv_fwd_and_tanh_bwd(cuda_lib, tc, C):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    x = _rand((3, 32, 32, C))
    w = torch.randn(5, 5, 3, C, device="cuda") * 0.05
    wb = w.to(torch.bfloat16) if tc else None
    if tc:
        w = wb.float()
    b = torch.randn(3, device="cuda") * 0.1
    out = K.c3_tconv(x, w, b, act="tanh", w_bf16=wb)
    ref = torch.tanh(Fn.deconv2d(x.float(), w, b))
    K.check_device_error()
    _close(out, ref, name="c3_tconv")
    aux = out
    cs = torch.zeros(3, device="cuda")
    out2 = K.c3_tconv(x, w, None, aux=aux, colsum=cs, w_bf16=wb)
    ref2 = Fn.deconv2d(x.float(), w) * (1 - aux.float() ** 2)
    _close(out2, ref2, name="c3_tconv_tanh_bwd")
    _close(cs, ref2.sum((0, 1, 2)), rel=2e-3, name="c3_tconv colsum")


@pytest.mark.parametrize("tc", [True, False])
@pytest.mark.parametrize("S,C", [(64, 64), (128, 128)])
def test_c3_wgrad(cuda_lib, tc, S, C):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    x3 = _rand((4, S, S, 3))
    y = _rand((4, S // 2, S // 2, C))
    dw = torch.zeros(5, 5, 3, C, device="cuda")
    K.c3_wgrad(x3, y, dw, tensor_cores=tc)
    w = torch.zeros(5, 5, 3, C, device="cuda", requires_grad=True)
    ref, = torch.autograd.grad(Fn.conv2d_same(x3.float(), w), w, y.float())
    K.check_device_error()
    _close(dw, ref, rel=2e-3, name="c3_wgrad")


@pytest.mark.parametrize("groups,rows,C,act", [(1, 1024, 512, "relu"), (2, 4096, 128, "lrelu"), (2, 512, 64, "lrelu")])
def test_bn_forward_backward(cuda_lib, groups, rows, C, act):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    x = _rand((groups * rows, C)) * 2 + 0.5
    gamma = 1 + 0.1 * torch.randn(C, device="cuda")
    beta = 0.1 * torch.randn(C, device="cuda")
    sums = torch.zeros(groups, 2, C, device="cuda")
    saved = torch.zeros(groups, 2, C, device="cuda")
    ema_m = torch.zeros(groups, C, device="cuda")
    ema_v = torch.zeros(groups, C, device="cuda")
    y = torch.empty_like(x)
    K.bn_stats(x, sums, groups)
    K.bn_apply(x, y, sums, gamma, beta, saved, ema_m, ema_v, groups, act)
    xr = x.float().view(groups, rows, C).requires_grad_(True)
    mean = xr.mean(1, keepdim=True)
    var = ((xr - mean) ** 2).mean(1, keepdim=True)
    yr = (xr - mean) * torch.rsqrt(var + 1e-5) * gamma + beta
    ar = torch.relu(yr) if act == "relu" else Fn.lrelu(yr)
    _close(y.view(groups, rows, C), ar, name="bn_apply")
    _close(ema_m, 0.1 * mean.squeeze(1), rel=2e-3, name="ema_mean")
    _close(ema_v, 0.1 * var.squeeze(1), rel=2e-3, name="ema_var")
    # backward with a third slab sharing the statistics of the last group
    slabs = groups + 1
    da = _rand((slabs * rows, C))
    red = torch.zeros(slabs, 2, C, device="cuda")
    dpre = torch.empty_like(da)
    dgamma = torch.zeros(C, device="cuda")
    dbeta = torch.zeros(C, device="cuda")
    K.bn_bwd_reduce(da, x, saved, gamma, beta, red, slabs, groups, act)
    K.bn_bwd_apply(da, x, dpre, saved, gamma, beta, red, dgamma, dbeta, slabs, groups, groups, act)
    daf = da.float().view(slabs, rows, C)
    gx, gg, gb = torch.autograd.grad(ar, (xr, gamma.requires_grad_(True), beta.requires_grad_(True)),
                                     daf[:groups], retain_graph=True) if False else (None, None, None)
    g2 = gamma.detach().clone().requires_grad_(True)
    b2 = beta.detach().clone().requires_grad_(True)
    xr2 = x.float().view(groups, rows, C).detach().requires_grad_(True)
    mean2 = xr2.mean(1, keepdim=True)
    var2 = ((xr2 - mean2) ** 2).mean(1, keepdim=True)
    yr2 = (xr2 - mean2) * torch.rsqrt(var2 + 1e-5) * g2 + b2
    ar2 = torch.relu(yr2) if act == "relu" else Fn.lrelu(yr2)
    gx, gg, gb = torch.autograd.grad(ar2, (xr2, g2, b2), daf[:groups], retain_graph=True)
    _close(dpre.view(slabs, rows, C)[:groups], gx, rel=2e-2, name="bn_bwd dx")
    _close(dgamma, gg, rel=1e-2, name="dgamma")
    _close(dbeta, gb, rel=1e-2, name="dbeta")
    # the extra slab: gradient through the last group's normalisation with its own upstream
    up = torch.zeros_like(daf[:groups])
    up[groups - 1] = daf[groups]
    gx3, = torch.autograd.grad(ar2, xr2, up)
    _close(dpre.view(slabs, rows, C)[groups], gx3[groups - 1], rel=2e-2, name="bn_bwd extra slab")


def test_act_bwd(cuda_lib):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    R, C = 512, 64
    a = _rand((2 * R, C))
    da = _rand((3 * R, C))
    dpre = torch.empty_like(da)
    dbias = torch.zeros(C, device="cuda")
    K.act_bwd(da, a, dpre, dbias, "lrelu", bias_rows=2 * R)
    a3 = torch.cat([a, a[R:]]).float()
    ref = da.float() * torch.where(a3 > 0, 1.0, 0.2)
    _close(dpre, ref, name="act_bwd")
    _close(dbias, ref[:2 * R].sum(0), rel=1e-2, name="act_bwd dbias")


def test_linear_and_head(cuda_lib):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    B, Kz, F = 64, 100, 8192
    z = torch.rand(B, Kz, device="cuda") * 2 - 1
    W = _rand((Kz, F), 0.02)
    b = torch.randn(F, device="cuda") * 0.01
    out = torch.empty(B, F, dtype=torch.bfloat16, device="cuda")
    K.lin_fwd(z, W, b, out)
    _close(out, z @ W.float() + b, name="lin_fwd")
    dy = _rand((B, F))
    dW = torch.empty(Kz, F, device="cuda")
    db = torch.empty(F, device="cuda")
    K.lin_wgrad(z, dy, dW, db)
    _close(dW, z.t() @ dy.float(), rel=2e-3, name="lin_wgrad")
    _close(db, dy.float().sum(0), rel=2e-3, name="lin_dbias")
    # discriminator head
    h = _rand((2 * B, F))
    w = torch.randn(F, device="cuda") * 0.02
    hb = torch.tensor([0.3], device="cuda")
    logits = torch.empty(2 * B, device="cuda")
    K.head_fwd(h, w, hb, logits)
    ref_logits = h.float() @ w + hb
    _close(logits, ref_logits, rel=2e-3, name="head_fwd")
    losses = torch.zeros(3, device="cuda")
    dlogit = torch.zeros(3 * B, device="cuda")
    probs = torch.zeros(2 * B, device="cuda")
    counters = torch.zeros(2, dtype=torch.int64, device="cuda")
    K.gan_loss(logits, losses, dlogit, probs, counters, B)
    lr_, lf_ = ref_logits[:B], ref_logits[B:]
    ref_losses = torch.stack([Fn.bce_with_logits_mean(lr_, 1.0), Fn.bce_with_logits_mean(lf_, 0.0),
                              Fn.bce_with_logits_mean(lf_, 1.0)])
    _close(losses, ref_losses, rel=1e-4, name="losses")
    assert counters.tolist() == [1, 1]
    ref_d = torch.cat([(torch.sigmoid(lr_) - 1) / B, torch.sigmoid(lf_) / B, (torch.sigmoid(lf_) - 1) / B])
    _close(dlogit, ref_d, rel=1e-3, name="dlogit")
    dh = torch.empty(3 * B, F, dtype=torch.bfloat16, device="cuda")
    dWh = torch.zeros(F, device="cuda")
    dbh = torch.zeros(1, device="cuda")
    K.head_bwd(h, w, dlogit, dh, dWh, dbh, B)
    _close(dh, dlogit[:, None] * w[None, :], name="head_bwd dh")
    _close(dWh, (dlogit[:2 * B, None] * h.float()).sum(0), rel=2e-3, name="head_bwd dW")
    _close(dbh, dlogit[:2 * B].sum().view(1), rel=2e-3, name="head_bwd db")
    # gan_loss + head_bwd in one launch: same outputs as the two kernels above
    losses2 = torch.zeros(3, device="cuda"); dlogit2 = torch.zeros(3 * B, device="cuda")
    probs2 = torch.zeros(2 * B, device="cuda"); counters2 = torch.zeros(2, dtype=torch.int64, device="cuda")
    dh2 = torch.empty_like(dh); dWh2 = torch.zeros(F, device="cuda"); dbh2 = torch.zeros(1, device="cuda")
    K.head_loss_bwd(h, w, logits, losses2, dlogit2, probs2, counters2, dh2, dWh2, dbh2, B)
    assert torch.equal(losses2, losses) and torch.equal(dlogit2, dlogit) and torch.equal(probs2, probs)
    assert counters2.tolist() == [1, 1]
    assert torch.equal(dh2, dh)
    _close(dWh2, dWh, rel=1e-5, name="fused head dW")
    _close(dbh2, dbh, rel=1e-5, name="fused head db")


def test_adam_matches_tf_form(cuda_lib):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    from distributed_tensorflow_for_dcgan_b200.optim.adam_tf import adam_tf_step_
    n = 100003
    p = torch.randn(n, device="cuda")
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    pb = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    pr, mr, vr = p.clone(), m.clone(), v.clone()
    counters = torch.zeros(2, dtype=torch.int64, device="cuda")
    for t in range(1, 4):
        g = torch.randn(n, device="cuda") * 0.01
        counters[0] = t
        K.adam(p, g, m, v, pb, counters)
        adam_tf_step_(pr, g, mr, vr, t)
    assert torch.allclose(p, pr, rtol=1e-5, atol=1e-6)
    assert torch.allclose(m, mr, rtol=1e-5, atol=1e-8)
    assert torch.equal(pb, p.to(torch.bfloat16))


def test_helpers(cuda_lib):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    src = _rand((25, 64, 128))
    dst = torch.empty(25, 128, 64, dtype=torch.bfloat16, device="cuda")
    K.transpose_taps(src, dst, 25, 64, 128)
    assert torch.equal(dst, src.transpose(1, 2).contiguous())
    x = _rand((5000, 3))
    out = torch.zeros(3, device="cuda")
    K.colsum(x, out)
    _close(out, x.float().sum(0), rel=1e-3, name="colsum")
    d = torch.rand(1000, dtype=torch.float64, device="cuda")
    o = torch.empty(1000, dtype=torch.bfloat16, device="cuda")
    K.decode_f64(d, o, 2.0, -1.0)
    assert torch.equal(o, (d.float() * 2 - 1).to(torch.bfloat16))
    img = (_rand((64, 8, 8, 3)).clamp(-1, 1))
    grid = K.grid_merge(img, 8)
    ref = ((img.float() + 1) / 2).view(8, 8, 8, 8, 3).permute(0, 2, 1, 3, 4).reshape(64, 64, 3)
    assert torch.allclose(grid, ref)
    zf = K.zero_fraction(torch.relu(_rand((4096,))))
    assert 0.3 < zf.item() < 0.7


def test_no_out_of_bounds_writes(cuda_lib):
    """compute-sanitizer is closed on the GPU pool, so out-of-bounds stores are caught with guard bands: every
    output lives inside a larger buffer pre-filled with a sentinel that must survive the kernel."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K

    def guarded(shape, dtype):
        n = 1
        for s in shape:
            n *= s
        pad = 4096
        buf = torch.full((n + 2 * pad,), 7.0 if dtype != torch.bfloat16 else 7.0, dtype=dtype, device="cuda")
        return buf, buf[pad:pad + n].view(shape), pad

    def intact(buf, pad):
        return bool((buf[:pad] == 7).all() and (buf[-pad:] == 7).all())

    x = _rand((3, 8, 8, 64))                       # 3*4*4 = 48 output rows: a partial 128-row tile
    w = _rand((25, 64, 64), 0.05)
    buf, out, pad = guarded((3, 4, 4, 64), torch.bfloat16)
    K.conv_fprop_kn(x, w, None, out=out)
    assert intact(buf, pad)
    buf, out, pad = guarded((3, 16, 16, 64), torch.bfloat16)
    K.conv_tconv(x, w, None, out=out)
    assert intact(buf, pad)
    buf, dw, pad = guarded((25, 64, 128), torch.float32)
    dw.zero_()
    K.conv_wgrad(_rand((4, 16, 16, 64)), _rand((4, 8, 8, 128)), dw)
    assert intact(buf, pad)
    x3 = _rand((2, 64, 64, 3))
    w3 = torch.randn(5, 5, 3, 64, device="cuda") * 0.05
    buf, out, pad = guarded((2, 32, 32, 64), torch.bfloat16)
    K.c3_fprop(x3, w3, None, out=out)
    assert intact(buf, pad)
    buf, out, pad = guarded((2, 64, 64, 3), torch.bfloat16)
    K.c3_tconv(_rand((2, 32, 32, 64)), w3, None, act="tanh", out=out)
    assert intact(buf, pad)
    buf, dw3, pad = guarded((5, 5, 3, 64), torch.float32)
    dw3.zero_()
    K.c3_wgrad(x3, _rand((2, 32, 32, 64)), dw3)
    assert intact(buf, pad)
    torch.cuda.synchronize()
    K.check_device_error()


def test_cluster_split_k_matches_single_cta(cuda_lib):
    """Small-M layers cut K across a thread-block cluster and reduce the partial accumulators through
    distributed shared memory; the result (and the fused batch-norm moments) must match the one-CTA-per-tile
    kernel up to f32 summation order."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    x = _rand((16, 8, 8, 256), 0.5)                 # 256 output rows of the strided conv -> 2 M tiles
    w_kn = _rand((25, 256, 128), 0.05)
    xt = _rand((16, 4, 4, 256), 0.5)                # transposed conv: 256 rows per parity class
    w_nk = _rand((25, 128, 256), 0.05)
    outs = {}
    try:
        for ks in (1, 2, 4):
            cuda_lib.dcgan_set_igemm_ksplit(ks)
            st = torch.zeros(2, 2, 128, device="cuda")
            o = torch.empty(16, 4, 4, 128, device="cuda", dtype=torch.bfloat16)
            K.conv_fprop_kn_stats(x, w_kn, None, o, st, 2, block_n=64)
            t = K.conv_tconv(xt, w_nk, None, block_n=64)
            K.check_device_error()
            outs[ks] = (o.float(), st.clone(), t.float())
    finally:
        cuda_lib.dcgan_set_igemm_ksplit(0)
    for ks in (2, 4):
        for a, b, name in zip(outs[ks], outs[1], ("fprop", "moments", "tconv")):
            _close(a, b, rel=1e-2, name=f"split-K {ks} {name}")
    # and against a plain fp32 PyTorch reference of the strided conv (TF SAME: pad 1 before, 2 after)
    xp = torch.nn.functional.pad(x.float().permute(0, 3, 1, 2), (1, 2, 1, 2))
    ref = torch.nn.functional.conv2d(xp, w_kn.float().view(5, 5, 256, 128).permute(3, 2, 0, 1), stride=2)
    _close(outs[4][0], ref.permute(0, 2, 3, 1), rel=2e-2, name="split-K 4 vs torch conv2d")


@pytest.mark.parametrize("transposed", [True, False])
def test_dgrad_epilogue_forms_bn_backward_reductions(cuda_lib, transposed):
    """The data-gradient conv's epilogue accumulates sum(dy) and sum(dy * xhat) of the batch-norm layer below
    (dy = da * act'(gamma * xhat + beta)) -- checked against plain fp32 PyTorch: conv(-transpose) for da, then the
    two reductions, per slab, with slab 2 reading the statistics of group 1 (the discriminator's g_loss slab)."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    slabs, groups = (3, 2) if transposed else (1, 1)
    B = 8
    N = slabs * B
    if transposed:   # da [N, 16, 16, 64] from x [N, 8, 8, 128]
        H, Cin, Cout, oH = 8, 128, 64, 16
        x = _rand((N, H, H, Cin), 0.5)
        w = _rand((25, Cout, Cin), 0.05)                                   # [tap][Cout][Cin]
        wt = w.float().view(5, 5, Cout, Cin).permute(3, 2, 0, 1)           # conv_transpose2d weight [Cin, Cout, kh, kw]
        full = torch.nn.functional.conv_transpose2d(x.float().permute(0, 3, 1, 2), wt, stride=2)
        da_ref = full[:, :, 1:1 + 2 * H, 1:1 + 2 * H].permute(0, 2, 3, 1).contiguous()
        act = "lrelu"
    else:            # da [N, 8, 8, 128] from x [N, 16, 16, 64]
        H, Cin, Cout, oH = 16, 64, 128, 8
        x = _rand((N, H, H, Cin), 0.5)
        w = _rand((25, Cin, Cout), 0.05)                                   # TensorFlow layout [tap][Cin][Cout]
        xp = torch.nn.functional.pad(x.float().permute(0, 3, 1, 2), (1, 2, 1, 2))
        da_ref = torch.nn.functional.conv2d(xp, w.float().view(5, 5, Cin, Cout).permute(3, 2, 0, 1), stride=2)
        da_ref = da_ref.permute(0, 2, 3, 1).contiguous()
        act = "relu"
    pre = _rand((groups * B, oH, oH, Cout))
    gamma = 1 + 0.1 * torch.randn(Cout, device="cuda")
    beta = 0.1 * torch.randn(Cout, device="cuda")
    saved = torch.empty(groups, 2, Cout, device="cuda")
    pf = pre.float().view(groups, -1, Cout)
    saved[:, 0] = pf.mean(1)
    saved[:, 1] = (pf.var(1, unbiased=False) + 1e-5).rsqrt()
    red = torch.zeros(slabs, 2, Cout, device="cuda")
    da = torch.empty(N, oH, oH, Cout, device="cuda", dtype=torch.bfloat16)
    K.conv_dgrad_bnred(x, w, da, pre, saved, gamma, beta, red, slabs, groups, act, transposed)
    K.check_device_error()
    _close(da, da_ref, name="data gradient")
    exp = torch.zeros_like(red)
    for g in range(slabs):
        sg = min(g, groups - 1)
        xh = (pf[sg] - saved[sg, 0]) * saved[sg, 1]
        y = gamma * xh + beta
        slope = (y > 0).float() if act == "relu" else torch.where(y > 0, 1.0, 0.2)
        dy = da_ref[g * B:(g + 1) * B].reshape(-1, Cout) * slope
        exp[g, 0] = dy.sum(0)
        exp[g, 1] = (dy * xh).sum(0)
    _close(red, exp, rel=5e-3, name="fused bn backward reductions")
    # and the stand-alone reduction kernel, fed the bf16 data gradient, agrees with the fused epilogue
    red2 = torch.zeros_like(red)
    K.bn_bwd_reduce(da.view(-1, Cout), pre.view(-1, Cout), saved, gamma, beta, red2, slabs, groups, act)
    _close(red, red2, rel=5e-3, name="fused vs stand-alone reduction")


def test_filter_gradient_is_reproducible(cuda_lib):
    """Split-K partial tiles are summed in slice order by the slice that arrives last, so two runs of the filter
    gradient agree bit for bit (the first-generation kernel accumulated with f32 atomics)."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    big, small = _rand((32, 16, 16, 64)), _rand((32, 8, 8, 128))
    outs = []
    for _ in range(3):
        dw = torch.zeros(25, 64, 128, device="cuda")
        K.conv_wgrad(big, small, dw, splits=5)
        outs.append(dw.clone())
    K.check_device_error()
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    x = _rand((16, 8, 8, 256), 0.5)
    w_kn = _rand((25, 256, 128), 0.05)
    try:
        cuda_lib.dcgan_set_igemm_ksplit(4)
        a = K.conv_fprop_kn(x, w_kn, None, block_n=64)
        b = K.conv_fprop_kn(x, w_kn, None, block_n=64)
    finally:
        cuda_lib.dcgan_set_igemm_ksplit(0)
    assert torch.equal(a, b)


@pytest.mark.parametrize("transposed", [True, False])
@pytest.mark.parametrize("ksplit", [0, 3])
def test_conv_with_batch_norm_in_its_launch(cuda_lib, transposed, ksplit):
    """conv + training-mode batch norm + activation in one launch (grid-wide barrier, then the elementwise tail)
    against fp32 PyTorch (conv, biased batch moments per group, affine, ReLU / LeakyReLU) and against the
    two-launch path (conv with moments, then bn_apply), including saved statistics and the EMA shadows."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    groups = 1 if transposed else 2
    N = 16
    if transposed:   # [N, 8, 8, 128] -> [N, 16, 16, 64]
        H, Cin, Cout, oH, act = 8, 128, 64, 16, "relu"
        x = _rand((N, H, H, Cin), 0.5)
        w = _rand((25, Cout, Cin), 0.05)
        wt = w.float().view(5, 5, Cout, Cin).permute(3, 2, 0, 1)
        full = torch.nn.functional.conv_transpose2d(x.float().permute(0, 3, 1, 2), wt, stride=2)
        ref = full[:, :, 1:1 + 2 * H, 1:1 + 2 * H].permute(0, 2, 3, 1).contiguous()
    else:            # [N, 32, 32, 64] -> [N, 16, 16, 128]
        H, Cin, Cout, oH, act = 32, 64, 128, 16, "lrelu"
        x = _rand((N, H, H, Cin), 0.5)
        w = _rand((25, Cin, Cout), 0.05)
        xp = torch.nn.functional.pad(x.float().permute(0, 3, 1, 2), (1, 2, 1, 2))
        ref = torch.nn.functional.conv2d(xp, w.float().view(5, 5, Cin, Cout).permute(3, 2, 0, 1), stride=2)
        ref = ref.permute(0, 2, 3, 1).contiguous()
    bias = 0.1 * torch.randn(Cout, device="cuda")
    gamma = 1 + 0.1 * torch.randn(Cout, device="cuda")
    beta = 0.1 * torch.randn(Cout, device="cuda")
    ref = ref + bias

    def buffers():
        return dict(pre=torch.empty(N, oH, oH, Cout, device="cuda", dtype=torch.bfloat16),
                    out=torch.empty(N, oH, oH, Cout, device="cuda", dtype=torch.bfloat16),
                    sums=torch.zeros(groups, 2, Cout, device="cuda"), saved=torch.zeros(groups, 2, Cout, device="cuda"),
                    em=torch.zeros(groups, Cout, device="cuda"), ev=torch.ones(groups, Cout, device="cuda"))
    a, b = buffers(), buffers()
    try:
        cuda_lib.dcgan_set_igemm_ksplit(ksplit)
        K.conv_bn_fwd(x, w, bias, a["pre"], a["out"], a["sums"], gamma, beta, a["saved"], a["em"], a["ev"], groups,
                      act, transposed, 1e-5, 0.9)
        (K.conv_tconv_stats if transposed else K.conv_fprop_kn_stats)(x, w, bias, b["pre"], b["sums"], groups)
    finally:
        cuda_lib.dcgan_set_igemm_ksplit(0)
    K.bn_apply(b["pre"].view(-1, Cout), b["out"].view(-1, Cout), b["sums"], gamma, beta, b["saved"], b["em"], b["ev"],
               groups, act, 1e-5, 0.9)
    K.check_device_error()
    # the two paths run the same arithmetic on the same bf16 pre-activations
    assert torch.equal(a["pre"], b["pre"])
    for k in ("sums", "saved", "em", "ev"):
        _close(a[k], b[k], rel=1e-5, name=f"one launch vs two: {k}")
    assert (a["out"].float() - b["out"].float()).abs().max().item() <= 2e-2
    # fp32 reference
    rf = ref.view(groups, -1, Cout)
    mean, var = rf.mean(1, keepdim=True), rf.var(1, unbiased=False, keepdim=True)
    y = gamma * (rf - mean) * (var + 1e-5).rsqrt() + beta
    y = torch.relu(y) if act == "relu" else torch.where(y > 0, y, 0.2 * y)
    _close(a["out"], y.view(N, oH, oH, Cout), rel=3e-2, name="conv + bn + act vs fp32")
    _close(a["saved"][:, 0], mean[:, 0], rel=2e-2, name="saved mean")
    _close(a["em"], 0.1 * mean[:, 0], rel=2e-2, name="ema mean")
    # the barrier words are zero again (workspace is reusable by the next launch) -- run it twice more
    for _ in range(2):
        c = buffers()
        K.conv_bn_fwd(x, w, bias, c["pre"], c["out"], c["sums"], gamma, beta, c["saved"], c["em"], c["ev"], groups,
                      act, transposed, 1e-5, 0.9)
        # (the moments are accumulated with f32 atomics, so repeats may differ in the last bf16 digit)
        assert (c["out"].float() - a["out"].float()).abs().max().item() <= 2e-2
    K.check_device_error()


@pytest.mark.parametrize("transposed", [True, False])
def test_dgrad_with_batch_norm_backward_in_its_launch(cuda_lib, transposed):
    """Data-gradient conv + the whole batch-norm backward pass of the layer below in one launch, against the
    three-launch path (conv, reduce, apply) and against autograd of fp32 batch norm."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    slabs, groups = (3, 2) if transposed else (1, 1)
    B = 8
    N = slabs * B
    if transposed:
        H, Cin, Cout, oH, act = 8, 128, 64, 16, "lrelu"
        x = _rand((N, H, H, Cin), 0.5)
        w = _rand((25, Cout, Cin), 0.05)
    else:
        H, Cin, Cout, oH, act = 16, 64, 128, 8, "relu"
        x = _rand((N, H, H, Cin), 0.5)
        w = _rand((25, Cin, Cout), 0.05)
    pre = _rand((groups * B, oH, oH, Cout))
    gamma = 1 + 0.1 * torch.randn(Cout, device="cuda")
    beta = 0.1 * torch.randn(Cout, device="cuda")
    saved = torch.empty(groups, 2, Cout, device="cuda")
    pf = pre.float().view(groups, -1, Cout)
    saved[:, 0] = pf.mean(1)
    saved[:, 1] = (pf.var(1, unbiased=False) + 1e-5).rsqrt()

    def bufs():
        return dict(da=torch.empty(N, oH, oH, Cout, device="cuda", dtype=torch.bfloat16),
                    dpre=torch.empty(N, oH, oH, Cout, device="cuda", dtype=torch.bfloat16),
                    red=torch.zeros(slabs, 2, Cout, device="cuda"), dg=torch.zeros(Cout, device="cuda"),
                    db=torch.zeros(Cout, device="cuda"))
    a, b = bufs(), bufs()
    K.conv_dgrad_bnbwd(x, w, a["da"], pre, saved, gamma, beta, a["red"], a["dpre"], a["dg"], a["db"], slabs, groups,
                       groups, act, transposed)
    K.conv_dgrad_bnred(x, w, b["da"], pre, saved, gamma, beta, b["red"], slabs, groups, act, transposed)
    K.bn_bwd_apply(b["da"].view(-1, Cout), pre.view(-1, Cout), b["dpre"].view(-1, Cout), saved, gamma, beta, b["red"],
                   b["dg"], b["db"], slabs, groups, groups, act)
    K.check_device_error()
    assert torch.equal(a["da"], b["da"])
    _close(a["red"], b["red"], rel=1e-4, name="reductions")
    _close(a["dg"], b["dg"], rel=1e-4, name="dgamma")
    _close(a["db"], b["db"], rel=1e-4, name="dbeta")
    assert (a["dpre"].float() - b["dpre"].float()).abs().max().item() <= 1e-2 * b["dpre"].float().abs().max().item()
    # autograd of fp32 batch norm fed the kernel's own (bf16) upstream gradient, slab by slab
    for g in range(slabs):
        sg = min(g, groups - 1)
        xin = pf[sg].clone().requires_grad_(True)
        mean, var = xin.mean(0), xin.var(0, unbiased=False)
        y = gamma * (xin - mean) * (var + 1e-5).rsqrt() + beta
        out = torch.relu(y) if act == "relu" else torch.where(y > 0, y, 0.2 * y)
        (gx,) = torch.autograd.grad(out, xin, a["da"][g * B:(g + 1) * B].float().reshape(-1, Cout))
        _close(a["dpre"][g * B:(g + 1) * B].reshape(-1, Cout), gx, rel=3e-2, name=f"dpre slab {g} vs autograd")


def test_one_launch_batch_norm_kernels(cuda_lib):
    """bn_forward_fused / bn_backward_fused (statistics, grid barrier, elementwise pass) against the two-launch
    kernels, for the shapes of g_bn0 / d_bn3 (512 channels, few rows) and g_bn3 (64 channels, many rows)."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    barrier = torch.zeros(2, dtype=torch.int32, device="cuda")
    for rows, C, groups, slabs, act in ((1024, 512, 1, 1, "relu"), (1024, 512, 2, 3, "lrelu"), (65536, 64, 1, 1, "relu")):
        x = _rand((groups * rows, C))
        gamma = 1 + 0.1 * torch.randn(C, device="cuda")
        beta = 0.1 * torch.randn(C, device="cuda")
        outs = []
        for fused in (True, False):
            y = torch.empty_like(x)
            sums, saved = torch.zeros(groups, 2, C, device="cuda"), torch.zeros(groups, 2, C, device="cuda")
            em, ev = torch.zeros(groups, C, device="cuda"), torch.ones(groups, C, device="cuda")
            if fused:
                K.bn_forward_fused(x, y, sums, gamma, beta, saved, em, ev, barrier, groups, act, 1e-5, 0.9)
            else:
                K.bn_stats(x, sums, groups)
                K.bn_apply(x, y, sums, gamma, beta, saved, em, ev, groups, act, 1e-5, 0.9)
            outs.append((y, sums, saved, em, ev))
        for u, v in zip(outs[0][1:], outs[1][1:]):
            _close(u, v, rel=1e-5, name="forward statistics")
        assert (outs[0][0].float() - outs[1][0].float()).abs().max().item() <= 2e-2
        saved = outs[1][2]
        da = _rand((slabs * rows, C))
        res = []
        for fused in (True, False):
            dpre = torch.empty_like(da)
            red = torch.zeros(slabs, 2, C, device="cuda")
            dg, db = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
            if fused:
                K.bn_backward_fused(da, x, dpre, saved, gamma, beta, red, dg, db, barrier, slabs, groups, groups, act)
            else:
                K.bn_bwd_reduce(da, x, saved, gamma, beta, red, slabs, groups, act)
                K.bn_bwd_apply(da, x, dpre, saved, gamma, beta, red, dg, db, slabs, groups, groups, act)
            res.append((dpre, red, dg, db))
        for u, v in zip(res[0][1:], res[1][1:]):
            _close(u, v, rel=1e-4, name="backward reductions")
        assert (res[0][0].float() - res[1][0].float()).abs().max().item() <= 1e-2 * res[1][0].float().abs().max().item()
    assert barrier.tolist() == [0, 0]
    K.check_device_error()


def test_tconv_with_activation_backward_in_its_epilogue(cuda_lib):
    """Data gradient into a bias + LeakyReLU layer without batch norm (d_h0): one launch against conv_tconv followed
    by act_bwd, three slabs with the g_loss slab reading the fake half's saved activations, bias gradient from the
    first two slabs only."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    B, H, Cin, Cout = 8, 8, 128, 64
    x = _rand((3 * B, H, H, Cin), 0.5)
    w = _rand((25, Cout, Cin), 0.05)
    act = _rand((2 * B, 2 * H, 2 * H, Cout))            # saved post-activation of the layer below
    out = torch.empty(3 * B, 2 * H, 2 * H, Cout, device="cuda", dtype=torch.bfloat16)
    dbias = torch.zeros(Cout, device="cuda")
    K.conv_tconv_actbwd(x, w, out, act, dbias, 3, 2, 2, "lrelu")
    da = K.conv_tconv(x, w, None)
    ref = torch.empty_like(out)
    dbias_ref = torch.zeros(Cout, device="cuda")
    K.act_bwd(da.view(-1, Cout), act.view(-1, Cout), ref.view(-1, Cout), dbias_ref, "lrelu",
              bias_rows=2 * B * 4 * H * H)
    K.check_device_error()
    # the fused path multiplies the f32 accumulator, the two-launch path the bf16-rounded data gradient
    assert (out.float() - ref.float()).abs().max().item() <= 1e-2 * ref.float().abs().max().item()
    _close(dbias, dbias_ref, rel=5e-3, name="bias gradient")
    slope = torch.where(torch.cat([act, act[B:]], 0).float() > 0, 1.0, 0.2)
    _close(out, da.float() * slope, rel=1e-2, name="dy vs da * lrelu'(a)")


def test_device_prefetcher_decodes_float64_records_on_the_gpu(cuda_lib, tmp_path):
    """Input pipeline with the decode on the device: raw float64 records -> pinned host -> H2D -> decode_f64_f32
    kernel on the prefetch stream.  Every row of every batch is the float32 cast of one of the records, bit exact."""
    import numpy as np
    from distributed_tensorflow_for_dcgan_b200.data import loader, native_reader, tfrecord
    rng = np.random.RandomState(1)
    imgs = rng.rand(48, 64, 64, 3) * 2 - 1
    d = tmp_path / "train"
    d.mkdir()
    for i in range(2):
        tfrecord.write_image_tfrecord(str(d / f"image_{i}.tfrecords"), imgs[i * 24:(i + 1) * 24])
    reader = native_reader.NativeShuffleReader(tfrecord.list_data_files(str(d)), capacity=40, min_after_dequeue=8,
                                               num_threads=2, seed=2, epochs=0, raw_f64=True)   # endless
    pf = loader.DevicePrefetcher(reader, 8, (64, 64, 3), "cuda", gpu_decode=True)
    got = []
    for _ in range(6):
        b = pf.next()
        assert b.dtype == torch.float32 and b.is_cuda
        got.append(b.cpu().numpy().copy())
        pf.release()
    pf.close()
    got = np.concatenate(got)
    ref = {a.tobytes() for a in imgs.astype(np.float32)}
    assert got.shape[0] == 48 and all(a.tobytes() in ref for a in got)   # every row is a record, cast bit-exactly
    assert len({a.tobytes() for a in got}) > 24                            # and they are not all the same few
    with pytest.raises(ValueError):
        loader.DevicePrefetcher(loader.SyntheticImages(8), 8, (64, 64, 3), "cuda", gpu_decode=True)
