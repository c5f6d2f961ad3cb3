"""Index arithmetic of the implicit-GEMM im2col gather (csrc/igemm_sm100.cu, modes F and T), emulated on the CPU.

`reference_*` is the straightforward per-k-block formula (tap and channel chunk from divisions, per-row
coordinates and bounds checks); `fast_*` is the strength-reduced form the kernel uses: per-row base offsets and
validity bit masks computed once, then one uniform offset and one bit index per k-block, advanced without
divisions.  Both must name the same source element (or zero fill) for every row, k-block and split-K slice."""
import itertools

import pytest


def parity_taps(ph, pw):
    return (3 if ph else 2), (3 if pw else 2)


def reference_seq(mode, R, kb_lo, kb_hi, gH, gW, gC, rH, rW, rows_total, ph, pw, chunk):
    """[(valid, element offset)] for row R over k-blocks [kb_lo, kb_hi)."""
    cpt = gC // 64
    stride = 1 if mode == "T" else 2
    out = []
    for kb in range(kb_lo, kb_hi):
        lt, c0 = kb // cpt, (kb % cpt) * 64
        if mode == "F":
            kh, kw = lt // 5, lt % 5
            dy, dx = kh - 1, kw - 1
        else:
            nkw = 3 if pw else 2
            ikh, ikw = lt // nkw, lt % nkw
            kh = 2 * ikh if ph else 2 * ikh + 1
            kw = 2 * ikw if pw else 2 * ikw + 1
            dy, dx = (ph + 1 - kh) // 2, (pw + 1 - kw) // 2
        if R >= rows_total:
            out.append((False, 0)); continue
        qw, qh, n = R % rW, (R // rW) % rH, R // (rW * rH)
        y, x = stride * qh + dy, stride * qw + dx
        ok = 0 <= y < gH and 0 <= x < gW
        out.append((ok, ((n * gH + y) * gW + x) * gC + c0 + chunk * 8 if ok else 0))
    return out


def fast_seq(mode, R, kb_lo, kb_hi, gH, gW, gC, rH, rW, rows_total, ph, pw, chunk):
    cpt = gC // 64
    stride = 1 if mode == "T" else 2
    # ---- once per thread and row ----
    if R < rows_total:
        qw, qh, n = R % rW, (R // rW) % rH, R // (rW * rH)
        y0, x0 = stride * qh, stride * qw
        base = ((n * gH + y0) * gW + x0) * gC + chunk * 8
        mask = 0
        for sy, sx in itertools.product(range(5), range(5)):
            if 0 <= y0 + sy - 1 < gH and 0 <= x0 + sx - 1 < gW:
                mask |= 1 << (sy * 5 + sx)
    else:
        base, mask = 0, 0
    # ---- once per CTA: position of the first k-block (the only division) ----
    lt, cc = kb_lo // cpt, kb_lo % cpt
    if mode == "F":
        a, b, nb = lt // 5, lt % 5, 5           # (kh, kw)
    else:
        nb = 3 if pw else 2
        a, b = lt // nb, lt % nb                # (ikh, ikw)
    out = []
    for _ in range(kb_lo, kb_hi):
        if mode == "F":
            dy, dx = a - 1, b - 1
        else:
            dy, dx = ph - a, pw - b
        delta = (dy * gW + dx) * gC + cc * 64
        bit = (dy + 1) * 5 + (dx + 1)
        ok = bool((mask >> bit) & 1)
        out.append((ok, base + delta if ok else 0))
        cc += 1                                  # advance without dividing
        if cc == cpt:
            cc = 0
            b += 1
            if b == nb:
                b, a = 0, a + 1
    return out


LAYERS = [  # mode, images, gathered H, gathered channels
    ("F", 4, 32, 64), ("F", 3, 16, 128), ("F", 2, 8, 256), ("F", 1, 4, 128),
    ("T", 4, 4, 512), ("T", 3, 8, 256), ("T", 2, 16, 128), ("T", 5, 4, 64),
]


@pytest.mark.parametrize("mode,NB,gH,gC", LAYERS)
def test_strength_reduced_gather_matches_reference(mode, NB, gH, gC):
    gW = gH
    rH = rW = gH if mode == "T" else gH // 2
    rows_total = NB * rH * rW
    cpt = gC // 64
    for ph, pw in ([(0, 0)] if mode == "F" else [(0, 0), (0, 1), (1, 0), (1, 1)]):
        num_kb = 25 * cpt if mode == "F" else parity_taps(ph, pw)[0] * parity_taps(ph, pw)[1] * cpt
        for ksplit in (1, 2, 4):
            if num_kb < ksplit:
                continue
            for krank in range(ksplit):
                lo, hi = krank * num_kb // ksplit, (krank + 1) * num_kb // ksplit
                rows = list(range(0, rows_total, max(1, rows_total // 37))) + [rows_total - 1, rows_total, rows_total + 5]
                for R in rows:
                    for chunk in (0, 5):
                        args = (mode, R, lo, hi, gH, gW, gC, rH, rW, rows_total, ph, pw, chunk)
                        assert fast_seq(*args) == reference_seq(*args), (ph, pw, ksplit, krank, R, chunk)
