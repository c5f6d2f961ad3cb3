"""CPU emulation of the index arithmetic of the TMA-fed implicit-GEMM kernels (csrc/igemm2_sm100.cu).

The kernels do no per-thread address generation: an activation tensor [N, H, W, C] is a 4-D tiled tensor map, a
128-row GEMM tile is one box {64 channels, bw pixels, bh rows, bn images} (``pixel_box``), the filter tap is a shift of
the box origin, TensorFlow's SAME padding is the TMA unit's out-of-bounds zero fill and the stride of the forward conv is
the map's element stride.  This file re-states that mapping in numpy -- ``tma_box`` is the TMA unit, the loops are the
producer warp of ``igemm2_kernel`` -- and checks the three modes against PyTorch on small shapes, so that the
coordinate formulas in the kernel source are pinned by a test that needs no GPU:

  MODE_F  strided conv      coords (c, tb - 1, 2 h0 + ta - 1, img0), element strides {1, 2, 2, 1}
  MODE_T  transposed conv   per output parity class (ph, pw): coords (c, pw - tb, h0 + ph - ta, img0), taps
                            kh = 2 ta (+1 if ph == 0), kw likewise; output pixel (2 qh + ph, 2 qw + pw)
  MODE_W  filter gradient   K = pixels: dense operand rows px0.., gathered operand box at (kw - 1, 2 hk + kh - 1, imgk)
(reference semantics: tf.nn.conv2d / conv2d_transpose, distriubted_model.py:183,200; their autodiff for MODE_W).
"""
import numpy as np
import pytest
import torch

F = torch.nn.functional


def pixel_box(rows, rH, rW):
    """igemm2_sm100.cu: pixel_box() -- `rows` consecutive pixels (x fastest, then y, then image) as a box."""
    if rW > rows or rows % rW:
        return None
    bw = rW
    if rH * rW >= rows:
        bh, bn = rows // rW, 1
        return (bw, bh, bn) if rH % bh == 0 else None
    bh, bn = rH, rows // (rH * rW)
    return (bw, bh, bn) if rows % (rH * rW) == 0 else None


def tma_box(x, c0, w0, h0, n0, box, stride):
    """The TMA unit: box {64, bw, bh, bn} of the [N, H, W, C] tensor at signed coordinates (c0, w0, h0, n0) with
    element strides {1, stride, stride, 1}; everything out of bounds reads as zero.  Returns [rows, 64] with
    row = (n * bh + h) * bw + w -- the order of the rows of a GEMM tile."""
    N, H, W, C = x.shape
    bw, bh, bn = box
    out = np.zeros((bn, bh, bw, 64), x.dtype)
    for n in range(bn):
        for h in range(bh):
            for w in range(bw):
                nn, hh, ww = n0 + n, h0 + h * stride, w0 + w * stride
                if 0 <= nn < N and 0 <= hh < H and 0 <= ww < W:
                    out[n, h, w] = x[nn, hh, ww, c0:c0 + 64]
    return out.reshape(bn * bh * bw, 64)


ROWS = 32   # the emulation uses 32-row tiles so that small shapes still span several tiles (the kernel uses 128)


@pytest.mark.parametrize("N,H,Cin,Cout", [(2, 8, 64, 64), (8, 4, 128, 64), (1, 16, 64, 128)])
def test_strided_conv_through_tma_boxes(N, H, Cin, Cout):
    rng = np.random.RandomState(0)
    x = rng.randn(N, H, H, Cin).astype(np.float32)
    w = rng.randn(25, Cin, Cout).astype(np.float32) * 0.1           # TensorFlow layout [tap][Cin][Cout]
    rH = rW = H // 2
    box = pixel_box(ROWS, rH, rW)
    assert box is not None
    rows_total = N * rH * rW
    out = np.zeros((rows_total, Cout), np.float32)
    for m0 in range(0, rows_total, ROWS):
        img0, h0 = m0 // (rH * rW), (m0 // rW) % rH                  # box origin; the box always starts at column 0
        acc = np.zeros((ROWS, Cout), np.float32)
        for ta in range(5):
            for tb in range(5):
                for cc in range(Cin // 64):
                    a = tma_box(x, cc * 64, tb - 1, 2 * h0 + ta - 1, img0, box, 2)
                    acc += a @ w[ta * 5 + tb, cc * 64:(cc + 1) * 64]
        out[m0:m0 + ROWS] = acc
    xp = F.pad(torch.from_numpy(x).permute(0, 3, 1, 2), (1, 2, 1, 2))      # SAME: 1 before, 2 after
    ref = F.conv2d(xp, torch.from_numpy(w).view(5, 5, Cin, Cout).permute(3, 2, 0, 1), stride=2)
    ref = ref.permute(0, 2, 3, 1).reshape(rows_total, Cout).numpy()
    assert np.abs(out - ref).max() <= 1e-3 * np.abs(ref).max()


@pytest.mark.parametrize("N,H,Cin,Cout", [(2, 4, 64, 64), (8, 2, 128, 64), (1, 8, 64, 64)])
def test_transposed_conv_through_tma_boxes(N, H, Cin, Cout):
    rng = np.random.RandomState(1)
    x = rng.randn(N, H, H, Cin).astype(np.float32)
    w = rng.randn(25, Cout, Cin).astype(np.float32) * 0.1           # [tap][Cout][Cin]
    rH = rW = H
    box = pixel_box(ROWS, rH, rW)
    assert box is not None
    rows_total = N * rH * rW
    out = np.zeros((N, 2 * H, 2 * H, Cout), np.float32)
    for ph in (1, 0):
        for pw in (1, 0):                                            # parity classes, longest K loop first
            na, nb = (3 if ph else 2), (3 if pw else 2)
            for m0 in range(0, rows_total, ROWS):
                img0, h0 = m0 // (rH * rW), (m0 // rW) % rH
                acc = np.zeros((ROWS, Cout), np.float32)
                for ta in range(na):
                    for tb in range(nb):
                        kh = 2 * ta if ph else 2 * ta + 1
                        kw = 2 * tb if pw else 2 * tb + 1
                        for cc in range(Cin // 64):
                            a = tma_box(x, cc * 64, pw - tb, h0 + ph - ta, img0, box, 1)
                            acc += a @ w[kh * 5 + kw, :, cc * 64:(cc + 1) * 64].T
                for r in range(ROWS):                                # the epilogue's scatter: output pixel 2q + p
                    R = m0 + r
                    qw, qh, n = R % rW, (R // rW) % rH, R // (rW * rH)
                    out[n, 2 * qh + ph, 2 * qw + pw] = acc[r]
    wt = torch.from_numpy(w).view(5, 5, Cout, Cin).permute(3, 2, 0, 1)     # conv_transpose2d weight [Cin, Cout, kh, kw]
    full = F.conv_transpose2d(torch.from_numpy(x).permute(0, 3, 1, 2), wt, stride=2)
    ref = full[:, :, 1:1 + 2 * H, 1:1 + 2 * H].permute(0, 2, 3, 1).numpy()  # TensorFlow's crop
    assert np.abs(out - ref).max() <= 1e-3 * np.abs(ref).max()


def test_filter_gradient_through_tma_boxes():
    rng = np.random.RandomState(2)
    N, Hs, Cg, Cd = 2, 4, 64, 128
    big = rng.randn(N, 2 * Hs, 2 * Hs, Cg).astype(np.float32)       # gathered at 2 q + k - 1
    small = rng.randn(N, Hs, Hs, Cd).astype(np.float32)             # dense operand, K = its pixels
    rows = N * Hs * Hs
    KB = 16                                                          # pixels per k-block in the emulation (kernel: 64)
    box = pixel_box(KB, Hs, Hs)
    assert box is not None
    dw = np.zeros((25, Cg, Cd), np.float32)
    s2 = small.reshape(rows, Cd)
    for tap in range(25):
        kh, kw = tap // 5, tap % 5
        for px0 in range(0, rows, KB):
            imgk, hk = px0 // (Hs * Hs), (px0 // Hs) % Hs
            g = tma_box(big, 0, kw - 1, 2 * hk + kh - 1, imgk, box, 2)      # [KB pixels, Cg]
            dw[tap] += g.T @ s2[px0:px0 + KB]                               # both operands MN-major, K = pixels
    x = torch.from_numpy(big).permute(0, 3, 1, 2)
    wt = torch.zeros(Cd, Cg, 5, 5, requires_grad=True)
    y = F.conv2d(F.pad(x, (1, 2, 1, 2)), wt, stride=2)
    (ref,) = torch.autograd.grad(y, wt, torch.from_numpy(small).permute(0, 3, 1, 2))
    ref = ref.permute(2, 3, 1, 0).reshape(25, Cg, Cd).numpy()               # HWIO = [tap][Cin][Cout]
    assert np.abs(dw - ref).max() <= 1e-3 * np.abs(ref).max()


def test_pixel_box_of_the_model_shapes():
    """Every layer of both model sizes decomposes into boxes within the TMA limits (extent <= 256 per dimension)."""
    for rows, grid in ((128, 4), (128, 8), (128, 16), (128, 32), (128, 64), (64, 4), (64, 8), (64, 16), (64, 32)):
        b = pixel_box(rows, grid, grid)
        assert b is not None and b[0] * b[1] * b[2] == rows
        assert b[0] * 2 <= 256 and b[1] * 2 <= 256 and b[2] <= 256
    assert pixel_box(128, 2, 2) == (2, 2, 32) and pixel_box(128, 64, 64) == (64, 2, 1)
    assert pixel_box(96, 8, 8) is None            # not a whole number of rows / images
