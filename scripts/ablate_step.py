"""Marginal cost of kernel families inside the graph-captured training step.

The step overlaps two streams, so a kernel's duration under a profiler says little about what it costs the step.
This script captures the step several times, each time leaving out the launches of one family of entry points
(the results of such a step are garbage; only its duration is used), and prints the step time and the delta to
the complete step.  CUDA events around graph replays, warm-up first, L2 flushed between replays.

usage: ablate_step.py [batch] [replays]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_for_dcgan_b200.config import ModelConfig
from distributed_tensorflow_for_dcgan_b200.engine import DCGANEngine
from distributed_tensorflow_for_dcgan_b200.ops import _lib, kernels as K

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
REPS = int(sys.argv[2]) if len(sys.argv) > 2 else 30

FAMILIES = [
    ("(complete step)", ()),
    ("bn_apply", ("dcgan_bn_apply",)),
    ("bn_bwd_apply", ("dcgan_bn_bwd_apply",)),
    ("bn_bwd_reduce+act_bwd+bn_stats", ("dcgan_bn_bwd_reduce", "dcgan_act_bwd", "dcgan_bn_stats")),
    ("all batch-norm kernels", ("dcgan_bn_apply", "dcgan_bn_bwd_apply", "dcgan_bn_bwd_reduce", "dcgan_bn_stats")),
    ("conv wgrad (side stream)", ("dcgan_conv_wgrad",)),
    ("c3 wgrad (side stream)", ("dcgan_c3_wgrad_tc",)),
    ("c3 fprop", ("dcgan_c3_fprop_tc",)),
    ("c3 tconv", ("dcgan_c3_tconv_tc",)),
    ("conv forward (stats variants)", ("dcgan_conv_fprop_kn_stats", "dcgan_conv_tconv_stats")),
    ("conv dgrad", ("dcgan_conv_tconv_bnred", "dcgan_conv_fprop_kn_bnred", "dcgan_conv_tconv", "dcgan_conv_fprop_kn")),
    ("adam", ("dcgan_adam",)),
    ("lin_fwd+lin_wgrad", ("dcgan_lin_fwd", "dcgan_lin_wgrad")),
    ("heads", ("dcgan_head_fwd", "dcgan_head_loss_bwd")),
    ("input cast", ("dcgan_f32_to_bf16",)),
]

orig_call = K.call
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def measure(skip):
    def call(name, *a):
        if name in skip:
            return
        orig_call(name, *a)
    K.call = call
    try:
        eng = DCGANEngine(ModelConfig(batch_size=B), seed=0)
        eng.real_f32.uniform_(-1, 1); eng.z.uniform_(-1, 1)
        eng.capture()
        for _ in range(5):
            eng.run_step()
        ts = []
        for _ in range(REPS):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); eng.run_step(); e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e) * 1e3)
        ts.sort()
        return ts[len(ts) // 2]
    finally:
        K.call = orig_call


base = None
for label, skip in FAMILIES:
    t = measure(set(skip))
    if base is None:
        base = t
    print(f"{t:8.1f} us   delta {base - t:7.1f} us   without {label}", flush=True)
