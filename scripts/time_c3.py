import os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_for_dcgan_b200.ops import _lib, kernels as K
lib = _lib.load()
lib.dcgan_set_c3_tuning.argtypes = [ctypes.c_int, ctypes.c_int]
x = (torch.randn(128, 64, 64, 3, device="cuda")).to(torch.bfloat16)
y = torch.randn(128, 32, 32, 64, device="cuda").to(torch.bfloat16)
w = torch.randn(5, 5, 3, 64, device="cuda") * 0.05
dw = torch.zeros(5, 5, 3, 64, device="cuda")
out = torch.empty(128, 32, 32, 64, dtype=torch.bfloat16, device="cuda")
def t(f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for blocks in (74, 148, 222, 296, 444, 592, 1024):
    lib.dcgan_set_c3_tuning(blocks, blocks)
    tf = t(lambda: K.c3_fprop(x, w, None, out=out))
    tw = t(lambda: K.c3_wgrad(x, y, dw))
    print(f"blocks {blocks:5d}: fprop_tc {tf:7.1f} us   wgrad_tc {tw:7.1f} us")
print("cuda-core: fprop", t(lambda: K.c3_fprop(x, w, None, out=out, tensor_cores=False)), "wgrad", t(lambda: K.c3_wgrad(x, y, dw, tensor_cores=False)))
xx = torch.randn(64, 32, 32, 64, device="cuda").to(torch.bfloat16)
oo = torch.empty(64, 64, 64, 3, dtype=torch.bfloat16, device="cuda")
print("c3_tconv", t(lambda: K.c3_tconv(xx, w, None, act="tanh", out=oo)))
wb = w.to(torch.bfloat16)
print("c3_tconv_tc", t(lambda: K.c3_tconv(xx, w, None, act="tanh", out=oo, w_bf16=wb)))
