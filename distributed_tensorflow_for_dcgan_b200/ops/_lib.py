"""ctypes binding of libdcgan_b200.so (C ABI in include/dcgan_b200.h).

The library is built in-tree by ``build_ext.py`` (nvcc, sm_100a).  On a machine with a GPU a missing
library is a hard error -- the CUDA path must never silently fall back to PyTorch ops.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libdcgan_b200.so")
_lib: Optional[ctypes.CDLL] = None

c_int, c_float, c_void_p, c_size_t = ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_size_t

# name -> argument kinds: p = device pointer, i = int, f = float, z = size_t, s = stream
_SIGNATURES = {
    "dcgan_conv_fprop": "ppppiiiiiiis",
    "dcgan_conv_fprop_kn": "ppppiiiiiiis",
    "dcgan_conv_tconv": "ppppiiiiiiis",
    "dcgan_conv_fprop_kn_stats": "pppppiiiiiiis",
    "dcgan_conv_tconv_stats": "pppppiiiiiiis",
    "dcgan_conv_wgrad": "pppiiiiiiis",
    "dcgan_conv_tconv_bnred": "ppppppppiiiiiiiiis",
    "dcgan_conv_fprop_kn_bnred": "ppppppppiiiiiiiiis",
    "dcgan_conv_tconv_bnbwd": "pppppppppppiiiiiiiiiis",
    "dcgan_conv_fprop_kn_bnbwd": "pppppppppppiiiiiiiiiis",
    "dcgan_conv_tconv_actbwd": "pppppiiiiiiiiiis",
    "dcgan_conv_fprop_kn_bn": "pppppppppppiiiiiiffiis",
    "dcgan_conv_tconv_bn": "pppppppppppiiiiiiffiis",
    "dcgan_bn_forward_fused": "pppppppppiiiffis",
    "dcgan_bn_backward_fused": "ppppppppppiiiiiis",
    "dcgan_c3_fprop": "ppppiiiiis",
    "dcgan_c3_tconv": "ppppppiiiiiis",
    "dcgan_c3_wgrad": "pppiiiis",
    "dcgan_c3_fprop_tc": "ppppiiiiis",
    "dcgan_c3_wgrad_tc": "pppiiiis",
    "dcgan_c3_tconv_tc": "ppppppiiiiiis",
    "dcgan_bn_stats": "ppiiis",
    "dcgan_bn_apply": "ppppppppiiiffiis",
    "dcgan_bn_bwd_reduce": "ppppppiiiiis",
    "dcgan_bn_bwd_apply": "pppppppppiiiiiis",
    "dcgan_act_bwd": "ppppiiiiis",
    "dcgan_lin_fwd": "ppppiiis",
    "dcgan_lin_wgrad": "ppppiiis",
    "dcgan_head_fwd": "ppppiis",
    "dcgan_gan_loss": "pppppifis",
    "dcgan_head_bwd": "ppppppiis",
    "dcgan_head_loss_bwd": "ppppppppppiifis",
    "dcgan_adam": "ppppppzfffffs",
    "dcgan_cast_bf16": "ppzs",
    "dcgan_transpose_taps": "ppiiis",
    "dcgan_colsum": "ppzis",
    "dcgan_decode_f64": "ppzffs",
    "dcgan_decode_f64_f32": "ppzffs",
    "dcgan_f32_to_bf16": "ppzffs",
    "dcgan_grid_merge": "ppiiiiis",
    "dcgan_count_zeros": "ppzs",
    "dcgan_comm_rs_adam_ag": "pppppppppiizzffffipppzzzis",
    "dcgan_comm_lin_fwd_gather": "ppppiiizzziis",
    "dcgan_comm_allreduce_small": "pppiipiiis",
    "dcgan_comm_async_apply": "ppppppppiizzziffffis",
    "dcgan_bn_forward_static": "pppppppiiiffis",
    "dcgan_copy_error_flag": "ps",
    "dcgan_zero_async": "pzs",
}
_KIND = {"p": c_void_p, "i": c_int, "f": c_float, "z": c_size_t, "s": c_void_p}


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise RuntimeError(
            f"{_LIB_PATH} is missing: run `python build_ext.py` (or __graft_entry__.build()) first. "
            "The CUDA path does not fall back to PyTorch ops.")
    lib = ctypes.CDLL(_LIB_PATH)
    for name, sig in _SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.argtypes = [_KIND[k] for k in sig]
        fn.restype = c_int
    lib.dcgan_last_device_error.argtypes = []
    lib.dcgan_last_device_error.restype = c_int
    lib.dcgan_set_mn_desc_variant.argtypes = [c_int]
    lib.dcgan_set_mn_desc_variant.restype = None
    lib.dcgan_set_workspace.argtypes = [c_void_p, c_size_t]
    lib.dcgan_set_workspace.restype = c_int
    for setter in ("dcgan_set_igemm_tuning", "dcgan_set_igemm_ksplit", "dcgan_set_wgrad_cta_target",
                   "dcgan_set_adam_blocks", "dcgan_set_igemm_max_ctas", "dcgan_set_igemm_stage_mode",
                   "dcgan_set_igemm_tile"):
        getattr(lib, setter).argtypes = [c_int]
        getattr(lib, setter).restype = None
    # profiling overrides (A/B runs of bench.py without a rebuild)
    if os.environ.get("DCGAN_IGEMM_TUNING"):
        lib.dcgan_set_igemm_tuning(int(os.environ["DCGAN_IGEMM_TUNING"]))
    if os.environ.get("DCGAN_IGEMM_KSPLIT"):
        lib.dcgan_set_igemm_ksplit(int(os.environ["DCGAN_IGEMM_KSPLIT"]))
    if os.environ.get("DCGAN_IGEMM_STAGE_MODE"):
        lib.dcgan_set_igemm_stage_mode(int(os.environ["DCGAN_IGEMM_STAGE_MODE"]))
    if os.environ.get("DCGAN_IGEMM_TILE"):
        lib.dcgan_set_igemm_tile(int(os.environ["DCGAN_IGEMM_TILE"]))
    if os.environ.get("DCGAN_IGEMM_MAX_CTAS"):
        lib.dcgan_set_igemm_max_ctas(int(os.environ["DCGAN_IGEMM_MAX_CTAS"]))
    if os.environ.get("DCGAN_WGRAD_CTAS"):
        lib.dcgan_set_wgrad_cta_target(int(os.environ["DCGAN_WGRAD_CTAS"]))
    if os.environ.get("DCGAN_ADAM_BLOCKS"):
        lib.dcgan_set_adam_blocks(int(os.environ["DCGAN_ADAM_BLOCKS"]))
    if os.environ.get("DCGAN_C3_TUNING"):  # "fprop_blocks,wgrad_blocks" of the 3-channel tensor-core kernels
        fb, wb = (int(v) for v in os.environ["DCGAN_C3_TUNING"].split(","))
        lib.dcgan_set_c3_tuning.argtypes = [c_int, c_int]
        lib.dcgan_set_c3_tuning.restype = None
        lib.dcgan_set_c3_tuning(fb, wb)
    _lib = lib
    return lib


def register(name: str, sig: str) -> None:
    """Late registration for entry points added by other modules (comm kernels)."""
    _SIGNATURES[name] = sig
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.argtypes = [_KIND[k] for k in sig]
        fn.restype = c_int


def ptr(t) -> Optional[int]:
    if t is None:
        return None
    if isinstance(t, int):
        return t
    return t.data_ptr()


def stream_handle() -> int:
    return torch.cuda.current_stream().cuda_stream


class KernelError(RuntimeError):
    pass


_ERR = {-1: "shape not supported", -2: "driver / tensor-map error", -3: "launch failed", -4: "size limit"}

# number of kernel launches issued through this binding (bench.py reports it as gpu_launches)
launch_count = 0


def call(name: str, *args) -> None:
    """Invoke an entry point; appends the current stream; raises on a non-zero return code."""
    global launch_count
    lib = load()
    fn = getattr(lib, name)
    rc = fn(*args, stream_handle())
    if rc != 0:
        raise KernelError(f"{name} failed: {_ERR.get(rc, rc)}")
    launch_count += 1


def set_workspace(t) -> None:
    """Split-K scratch for the implicit-GEMM launches of this host thread (``None`` = the library default)."""
    rc = load().dcgan_set_workspace(None if t is None else t.data_ptr(), 0 if t is None else t.numel() * t.element_size())
    if rc != 0:
        raise KernelError(f"dcgan_set_workspace failed: {_ERR.get(rc, rc)}")


def last_device_error() -> int:
    return load().dcgan_last_device_error()
