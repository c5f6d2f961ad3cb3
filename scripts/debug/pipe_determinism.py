import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from distributed_tensorflow_for_dcgan_b200.config import ModelConfig
from distributed_tensorflow_for_dcgan_b200.engine import DCGANEngine
cfg = ModelConfig(batch_size=16)
g = torch.Generator().manual_seed(5)
reals = [(torch.rand(16, 64, 64, 3, generator=g) * 2 - 1).pin_memory() for _ in range(3)]
zs = [(torch.rand(16, cfg.z_dim, generator=g) * 2 - 1).pin_memory() for _ in range(3)]
def mk(graph=True):
    e = DCGANEngine(cfg, seed=2)
    if graph: e.capture()
    return e
a, b = mk(), mk()
print("init equal", torch.equal(a.arena.master, b.arena.master))
for name, e in (("sync-a", a), ("sync-b", b)):
    print(name, [e.train_step(r, z)[:3] for r, z in zip(reals, zs)])
c = mk()
out, prev = [], None
for r, z in zip(reals, zs):
    t = c.submit(r, z)
    if prev is not None: out.append(c.result(prev)[:3])
    prev = t
out.append(c.result(prev)[:3])
print("pipe-c", out)
d = mk(False)
print("eager-d", [d.train_step(r, z)[:3] for r, z in zip(reals, zs)])
d2 = mk(False)
d2.real_f32.copy_(reals[0]); d2.z.copy_(zs[0]); torch.cuda.synchronize(); d2.run_step(); torch.cuda.synchronize()
print("eager-direct", d2.losses.tolist())
