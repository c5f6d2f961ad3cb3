"""Prints per-variable gradient error of one engine step against the fp32 oracle (GPU)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_engine_gpu import _oracle_step
from distributed_tensorflow_for_dcgan_b200.config import ModelConfig
from distributed_tensorflow_for_dcgan_b200.engine import DCGANEngine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
cfg = ModelConfig(batch_size=B)
eng = DCGANEngine(cfg, seed=3)
torch.manual_seed(0)
real = torch.rand(B, 64, 64, 3, device="cuda") * 2 - 1
z = torch.rand(B, cfg.z_dim, device="cuda") * 2 - 1
p0 = {k: (v.detach().to(torch.bfloat16).float() if (k.endswith("/w") or k == "g_h0_lin/Matrix") else v.detach().clone())
      for k, v in eng.params.items()}
eng.real_f32.copy_(real); eng.z.copy_(z)
eng.step_body(); torch.cuda.synchronize()
losses, grads, G = _oracle_step(p0, cfg, real.to(torch.bfloat16).float(), z)
print("losses", eng.losses.tolist(), losses)
for name, ref in grads.items():
    gv = eng.grad_views[name]
    err = (gv - ref).abs().max().item(); den = ref.abs().max().item()
    cos = torch.nn.functional.cosine_similarity(gv.flatten(), ref.flatten(), dim=0).item()
    print(f"{name:24s} err/max={err/max(den,1e-30):8.4f} max={den:10.4g} cos={cos:.5f}")
