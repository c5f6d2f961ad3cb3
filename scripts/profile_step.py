"""Runs N un-graphed engine steps so that ncu can list / profile every kernel of one step.
usage: profile_step.py [steps] [batch]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_for_dcgan_b200.config import ModelConfig
from distributed_tensorflow_for_dcgan_b200.engine import DCGANEngine
from distributed_tensorflow_for_dcgan_b200.ops import _lib, kernels as K

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
cfg = ModelConfig(batch_size=B)
eng = DCGANEngine(cfg, seed=0)
eng.real_f32.uniform_(-1, 1); eng.z.uniform_(-1, 1)
torch.cuda.synchronize()
print("init launches", _lib.launch_count)
for i in range(steps):
    c0 = _lib.launch_count
    last = i == steps - 1
    if last:
        torch.cuda.profiler.start()  # ncu --profile-from-start off: exactly the last step is profiled
    eng.step_body()
    torch.cuda.synchronize()
    if last:
        torch.cuda.profiler.stop()
    print("step", i, "launches", _lib.launch_count - c0, eng.losses.tolist())
K.check_device_error()
