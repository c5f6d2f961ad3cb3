"""In-situ timing of the phases / individual kernels of one training step (CUDA events, warm L2, no profiler)."""
import os, sys, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_for_dcgan_b200.config import ModelConfig
from distributed_tensorflow_for_dcgan_b200.engine import DCGANEngine
from distributed_tensorflow_for_dcgan_b200.ops import _lib, kernels as K

cfg = ModelConfig(batch_size=64)
eng = DCGANEngine(cfg, seed=0)
eng.real_f32.uniform_(-1, 1); eng.z.uniform_(-1, 1)
for _ in range(3): eng.step_body()
torch.cuda.synchronize()

# per-kernel in-situ timing: wrap _lib.call with events
events = []
orig_call = _lib.call
def timed_call(name, *a):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); orig_call(name, *a); e.record()
    events.append((name, s, e))
K.call = timed_call
reps = 5
for _ in range(reps):
    eng.step_body()
torch.cuda.synchronize()
agg = collections.OrderedDict()
for name, s, e in events:
    agg.setdefault(name, [0.0, 0])
    agg[name][0] += s.elapsed_time(e) * 1e3 / reps
    agg[name][1] += 1
tot = sum(v[0] for v in agg.values())
for name, (us, n) in sorted(agg.items(), key=lambda x: -x[1][0]):
    print(f"{us:8.1f} us  x{n // reps:2d}  {name}")
print(f"sum of kernel spans {tot:.1f} us (event pairs add ~2 us each when kernels are short)")
K.call = orig_call
# phases
def t(f):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record(); f(); e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) * 1e3
eng.zero_region.zero_()
print("G_forward", t(lambda: eng.generator_forward(True)))
print("D_forward", t(eng.discriminator_forward))
K.gan_loss(eng.logits, eng.losses, eng.dlogit, eng.probs, eng.counters, eng.B, 1.0, True)
print("D_backward", t(eng.discriminator_backward))
print("G_backward", t(eng.generator_backward))
print("optimizer", t(lambda: eng.optimizer_step("both")))
