"""Run the forward pass twice from the same state and report where the two runs first differ."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from distributed_tensorflow_for_dcgan_b200.config import ModelConfig
from distributed_tensorflow_for_dcgan_b200.engine import DCGANEngine
B = int(os.environ.get("B", "16"))
cfg = ModelConfig(batch_size=B)
e = DCGANEngine(cfg, seed=2)
g = torch.Generator().manual_seed(5)
e.real_f32.copy_(torch.rand(B, 64, 64, 3, generator=g) * 2 - 1); e.z.copy_(torch.rand(B, cfg.z_dim, generator=g) * 2 - 1)
state = e.snapshot_state()
runs = []
for r in range(2):
    e.restore_state(state)
    # what step_body does before the forward pass
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    e.grads.zero_(); e._scratch.zero_()
    e.generator_forward(True)
    snap = {}
    for i in range(4):
        snap[f"g_pre{i}"] = e.g_pre[i].float().clone(); snap[f"g_act{i}"] = e.g_act[i].float().clone()
        snap[f"g_bn{i}.sums"] = e.bn[f"g_bn{i}"]["sums"].clone()
    snap["fake"] = e.d_in[B:].float().clone()
    runs.append(snap)
    torch.cuda.synchronize()
for k in runs[0]:
    a, b = runs[0][k], runs[1][k]
    d = (a - b).abs().max().item()
    print(f"{k:12s} maxdiff {d:.3e}  rel {d / (a.abs().max().item() + 1e-30):.3e}  mismatching {(a != b).float().mean().item():.3e}")
