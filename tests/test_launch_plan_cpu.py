"""Host-side launch plan of the implicit-GEMM convs (tile width, cluster split-K factor, grid): no GPU needed.

The plan is what ``dcgan_conv_fprop*`` / ``dcgan_conv_tconv*`` use on the device (csrc/igemm_sm100.cu:
pick_block_n / pick_ksplit); the layer shapes are those of SURVEY.md 2.4 (a)-(c) at the reference batch of 64
(D runs on [real; fake] = 128 images forward and 3 slabs = 192 images backward)."""
import ctypes
import os

import pytest

from distributed_tensorflow_for_dcgan_b200.ops import _lib


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_lib.lib_path()):
        pytest.skip("libdcgan_b200.so not built (python build_ext.py)")
    try:
        return ctypes.CDLL(_lib.lib_path())
    except OSError as e:  # no CUDA runtime on this machine
        pytest.skip(f"cannot load the library here: {e}")


def plan(lib, transposed, NB, H, Cin, Cout):
    bn, ks, fit = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    grid = (ctypes.c_int * 3)()
    rc = lib.dcgan_debug_plan_conv(transposed, NB, H, H, Cin, Cout, ctypes.byref(bn), ctypes.byref(ks), grid,
                                   ctypes.byref(fit))
    assert rc == 0
    return bn.value, ks.value, tuple(grid), bool(fit.value)


LAYERS = {  # name: (transposed, images, input H, Cin, Cout)
    "d_h1 forward": (0, 128, 32, 64, 128), "d_h2 forward": (0, 128, 16, 128, 256), "d_h3 forward": (0, 128, 8, 256, 512),
    "g_h1 data grad": (0, 64, 8, 256, 512), "g_h2 data grad": (0, 64, 16, 128, 256), "g_h3 data grad": (0, 64, 32, 64, 128),
    "g_h1 forward": (1, 64, 4, 512, 256), "g_h2 forward": (1, 64, 8, 256, 128), "g_h3 forward": (1, 64, 16, 128, 64),
    "d_h3 data grad": (1, 192, 4, 512, 256), "d_h2 data grad": (1, 192, 8, 256, 128), "d_h1 data grad": (1, 192, 16, 128, 64),
}


def test_plan_invariants(lib):
    for name, (t, nb, h, cin, cout) in LAYERS.items():
        bn, ks, grid, fit = plan(lib, t, nb, h, cin, cout)
        assert bn in (64, 128, 256) and cout % bn == 0, name
        assert ks in (1, 2, 4) and fit, name                      # partial tiles fit the leader's stage ring
        assert grid[0] % ks == 0, name                            # clusters of ks CTAs tile grid.x exactly
        k_blocks = (4 if t else 25) * (cin // 64)                 # shortest K loop (2x2-tap parity class / 25 taps)
        assert k_blocks // ks >= (8 if t else 16) or ks == 1, name
        slots = 148 * (1 if bn == 256 else 2)
        assert ks == 1 or grid[0] * grid[1] * grid[2] <= slots, name  # split-K never spills into a second wave


def test_small_m_layers_are_split(lib):
    """The layers that leave most SMs idle with one CTA per tile (profiles/igemm_splitk_r1.log)."""
    assert plan(lib, *LAYERS["g_h1 data grad"])[:2] == (64, 4)
    assert plan(lib, *LAYERS["d_h3 forward"])[:2] == (128, 2)
    assert plan(lib, *LAYERS["g_h1 forward"])[:2] == (128, 2)
    # large-M layers already fill the machine
    assert plan(lib, *LAYERS["d_h1 data grad"])[1] == 1
    assert plan(lib, *LAYERS["d_h1 forward"])[1] == 1


def test_rejects_unsupported_channels(lib):
    bn, ks, fit = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    grid = (ctypes.c_int * 3)()
    assert lib.dcgan_debug_plan_conv(0, 8, 16, 16, 48, 64, ctypes.byref(bn), ctypes.byref(ks), grid,
                                     ctypes.byref(fit)) < 0


def test_header_declares_every_exported_entry_point():
    """include/dcgan_b200.h is the C ABI: every dcgan_* symbol the library exports is declared there, and
    everything declared is exported."""
    import re
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.exists(_lib.lib_path()):
        pytest.skip("library not built")
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.lib_path()], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if l.split() and l.split()[-1].startswith("dcgan_")}
    header = open(os.path.join(root, "include", "dcgan_b200.h")).read()
    declared = set(re.findall(r"\b(dcgan_[a-z0-9_]+)\s*\(", header))
    assert exported == declared, (sorted(exported - declared), sorted(declared - exported))
