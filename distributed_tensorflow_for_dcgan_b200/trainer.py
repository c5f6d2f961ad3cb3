"""Trainers: one interface, two compute paths.

* ``CPUTrainer``    -- PyTorch oracle model + autograd (fp32); the plumbing path (gloo, parameter
  servers, BASELINE.json config 1) and the reference against which the engine is tested.
* ``EngineTrainer`` -- the B200 engine (hand-written sm_100a kernels, CUDA-graph replay).

Both implement the fetches of the reference's step (``image_train.py:156-158``), the sample-time
evaluation (``:181-184``), the summaries (``:86-89,98-101,114-118``; ``distriubted_model.py:75-80``)
and a TF-named ``state_dict``.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np
import torch

from .config import ModelConfig
from .models import dcgan, manifest
from .ops import functional as Fn
from .optim.adam_tf import AdamTF
from .utils.summary import Summary


class CPUTrainer:
    def __init__(self, cfg: ModelConfig, lr=2e-4, beta1=0.5, seed=0, update_mode="simultaneous", device="cpu",
                 ps_client=None, allreduce_world: int = 1, sync_bn: bool = False):
        self.cfg, self.device, self.update_mode = cfg, torch.device(device), update_mode
        self.params = dcgan.init_params(cfg, seed=seed, device=self.device)
        self.ema = dcgan.init_ema(cfg, device=self.device)
        self.d_names = [n for n in self.params if manifest.is_d_var(n)]
        self.g_names = [n for n in self.params if manifest.is_g_var(n)]
        self.opt_d = AdamTF(self.params, self.d_names, lr, beta1)
        self.opt_g = AdamTF(self.params, self.g_names, lr, beta1)
        self.global_step = 0
        self.ps = ps_client
        self.allreduce_world = allreduce_world
        self.sync_bn = sync_bn
        self.last_acts: Dict[str, torch.Tensor] = {}
        self.last_G: Optional[torch.Tensor] = None
        self.last_D = (None, None)

    # ---- one pass of the reference graph ------------------------------------------------------------------
    def _forward(self, real, z, p, record=False, update_ema=True):
        acts = {} if record else None

        def bn_hook(scope, mean, var):
            if update_ema:
                m, v = manifest.ema_names(scope)
                Fn.ema_update_(self.ema[m], mean, self.cfg.bn_momentum)
                Fn.ema_update_(self.ema[v], var, self.cfg.bn_momentum)
        act_hook = (lambda name, t: acts.__setitem__(name, t.detach())) if record else None
        G = dcgan.generator(z, p, self.cfg, bn_hook, act_hook)
        D, D_logits = dcgan.discriminator(real, p, self.cfg, False, bn_hook, act_hook)
        D_, D_logits_ = dcgan.discriminator(G, p, self.cfg, True, bn_hook, act_hook)
        d_real, d_fake, g_loss = dcgan.gan_losses(D_logits, D_logits_)
        if record:
            self.last_acts, self.last_G, self.last_D = acts, G.detach(), (D.detach(), D_.detach())
        return G, d_real, d_fake, g_loss

    def _grads(self, real, z, record=False):
        p = {k: v.requires_grad_(True) for k, v in self.params.items()}
        _, d_real, d_fake, g_loss = self._forward(real, z, p, record)
        dg = torch.autograd.grad(d_real + d_fake, [p[n] for n in self.d_names], retain_graph=True)
        gg = torch.autograd.grad(g_loss, [p[n] for n in self.g_names])
        for v in self.params.values():
            v.requires_grad_(False)
        grads = dict(zip(self.d_names, dg))
        grads.update(dict(zip(self.g_names, gg)))
        return (d_real.item(), d_fake.item(), g_loss.item()), grads

    def train_step(self, real, z, record: bool = False) -> Tuple[float, float, float, int]:
        real = torch.as_tensor(real, dtype=torch.float32, device=self.device)
        z = torch.as_tensor(z, dtype=torch.float32, device=self.device)
        if self.update_mode == "alternating" and self.ps is None:
            return self._alternating_step(real, z, record)
        losses, grads = self._grads(real, z, record)
        if self.ps is not None:                      # parameter servers apply Adam and own global_step
            self.global_step = self.ps.push_pull(grads)
        else:
            if self.allreduce_world > 1:             # ps-less synchronous replicas over gloo
                import torch.distributed as dist
                for n in grads:
                    g = grads[n].contiguous()
                    dist.all_reduce(g)
                    grads[n] = g / self.allreduce_world
            self.opt_d.step(grads)
            self.opt_g.step(grads)
            self.global_step += 1                    # g_optim increments global_step (image_train.py:112)
        return losses[0], losses[1], losses[2], self.global_step

    def _alternating_step(self, real, z, record):
        p = {k: v.requires_grad_(True) for k, v in self.params.items()}
        _, d_real, d_fake, _ = self._forward(real, z, p, record)
        dg = torch.autograd.grad(d_real + d_fake, [p[n] for n in self.d_names])
        for v in self.params.values():
            v.requires_grad_(False)
        self.opt_d.step(dict(zip(self.d_names, dg)))
        p = {k: v.requires_grad_(True) for k, v in self.params.items()}
        _, _, _, g_loss = self._forward(real, z, p, False, update_ema=False)
        gg = torch.autograd.grad(g_loss, [p[n] for n in self.g_names])
        for v in self.params.values():
            v.requires_grad_(False)
        self.opt_g.step(dict(zip(self.g_names, gg)))
        self.global_step += 1
        return d_real.item(), d_fake.item(), g_loss.item(), self.global_step

    # ---- sampling (image_train.py:179-192) ----------------------------------------------------------------
    @torch.no_grad()
    def evaluate_losses(self, real, z) -> Tuple[float, float]:
        """d_loss / g_loss through the *training-mode* graph, EMA update firing as a side effect, exactly
        as the reference's sample-time ``sess.run([sampler, d_loss, g_loss])`` does."""
        real = torch.as_tensor(real, dtype=torch.float32, device=self.device)
        z = torch.as_tensor(z, dtype=torch.float32, device=self.device)
        _, d_real, d_fake, g_loss = self._forward(real, z, self.params)
        return (d_real + d_fake).item(), g_loss.item()

    @torch.no_grad()
    def sample(self, z) -> np.ndarray:
        z = torch.as_tensor(z, dtype=torch.float32, device=self.device)
        return dcgan.sampler(z, self.params, self.ema, self.cfg).cpu().numpy()

    def sample_grid(self, z, cols: int = 8) -> np.ndarray:
        from .utils.sample import inverse_transform, merge
        imgs = self.sample(z)
        return merge(inverse_transform(imgs), [(imgs.shape[0] + cols - 1) // cols, cols])

    # ---- summaries ------------------------------------------------------------------------------------------
    def summaries(self, real, z, losses) -> bytes:
        """The merged ``summary_op``: needs a recorded forward pass (``train_step(.., record=True)``)."""
        s = Summary()
        s.histogram("z", np.asarray(z))
        if self.last_G is not None:
            s.image("G", self.last_G.cpu().numpy())
            s.histogram("d", self.last_D[0].cpu().numpy())
            s.histogram("d_", self.last_D[1].cpu().numpy())
        d_real, d_fake, g_loss = losses
        s.scalar("d_loss_real", d_real); s.scalar("d_loss_fake", d_fake)
        s.scalar("g_loss", g_loss); s.scalar("d_loss", d_real + d_fake)
        for name, t in self.last_acts.items():
            s.histogram(name + "/activations", t.cpu().numpy())
            s.scalar(name + "/sparsity", float(Fn.zero_fraction(t)))
        for name, t in self.params.items():
            s.histogram(name, t.detach().cpu().numpy())
        return s.serialize()

    # ---- state ---------------------------------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, torch.Tensor]:
        sd = {n: t.detach().clone() for n, t in self.params.items()}
        sd.update({n: t.clone() for n, t in self.ema.items()})
        sd[manifest.GLOBAL_STEP_NAME] = torch.tensor(self.global_step, dtype=torch.int32)
        if self.ps is not None:
            m, v = self.ps.get_slots()
            for n in self.params:
                a, b = manifest.adam_slot_names(n)
                sd[a], sd[b] = m[n], v[n]
            sd["adam_t"] = torch.tensor(self.ps.t, dtype=torch.int64)
        else:
            sd.update(self.opt_d.state_dict())
            sd.update(self.opt_g.state_dict())
        return sd

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        with torch.no_grad():
            for n, t in self.params.items():
                t.copy_(sd[n])
            for n, t in self.ema.items():
                if n in sd:
                    t.copy_(sd[n])
        self.opt_d.load_state_dict(sd)
        self.opt_g.load_state_dict(sd)
        if manifest.GLOBAL_STEP_NAME in sd:
            self.global_step = int(sd[manifest.GLOBAL_STEP_NAME])

    def adam_slots(self):
        m = {**self.opt_d.m, **self.opt_g.m}
        v = {**self.opt_d.v, **self.opt_g.v}
        return m, v


class EngineTrainer:
    """The B200 path behind the same interface."""

    def __init__(self, cfg: ModelConfig, lr=2e-4, beta1=0.5, seed=0, update_mode="simultaneous", device="cuda",
                 comm=None, cuda_graph=True):
        from .engine import DCGANEngine
        self.cfg = cfg
        self.engine = DCGANEngine(cfg, device=device, seed=seed, lr=lr, beta1=beta1, update_mode=update_mode,
                                  comm=comm)
        if comm is not None:
            comm.attach(self.engine)
        self._want_graph = cuda_graph
        self.device = self.engine.device

    @property
    def global_step(self) -> int:
        comm = self.engine.comm
        if comm is not None and hasattr(comm, "global_step"):   # async PS: applies of *all* workers
            return comm.global_step()
        return int(self.engine.counters[1].item())

    @property
    def params(self):
        return self.engine.params

    def train_step(self, real, z, record: bool = False):
        e = self.engine
        if self._want_graph and e._graph is None:
            e.capture()
        real = torch.as_tensor(real)
        z = torch.as_tensor(z)
        d_real, d_fake, g_loss, step = e.train_step(real, z)
        if e.comm is not None and hasattr(e.comm, "global_step"):
            step = e.comm.global_step()
        return d_real, d_fake, g_loss, step

    def evaluate_losses(self, real, z):
        """Training-mode forward only (EMA side effect included), no parameter update."""
        from .ops import kernels as K
        e = self.engine
        e.real_f32.copy_(torch.as_tensor(real)); e.z.copy_(torch.as_tensor(z))
        e.zero_region.zero_()
        K.f32_to_bf16(e.real_f32, e.d_in[:e.B])
        # only the chief evaluates (image_train.py:179-184): batch statistics stay local, no collective is entered
        e._local_bn_only = True
        try:
            e.generator_forward(train=True)
            e.discriminator_forward()
        finally:
            e._local_bn_only = False
        K.gan_loss(e.logits, e.losses, None, e.probs, None, e.B, 1.0, False)
        l = e.losses.tolist()
        return l[0] + l[1], l[2]

    def sample(self, z) -> np.ndarray:
        return self.engine.sample(torch.as_tensor(z).to(self.device)).float().cpu().numpy()

    def sample_grid(self, z, cols: int = 8) -> np.ndarray:
        """inverse_transform + merge on the device; only the finished grid crosses PCIe."""
        from .ops import kernels as K
        imgs = self.engine.sample(torch.as_tensor(z).to(self.device))
        return K.grid_merge(imgs, cols).cpu().numpy()

    def summaries(self, real, z, losses) -> bytes:
        from .ops import kernels as K
        e = self.engine
        B = e.B
        s = Summary()
        s.histogram("z", np.asarray(torch.as_tensor(z).cpu()))
        s.image("G", e.d_in[B:].float().cpu().numpy())
        s.histogram("d", e.probs[:B].cpu().numpy()); s.histogram("d_", e.probs[B:].cpu().numpy())
        d_real, d_fake, g_loss = losses
        s.scalar("d_loss_real", d_real); s.scalar("d_loss_fake", d_fake)
        s.scalar("g_loss", g_loss); s.scalar("d_loss", d_real + d_fake)
        acts = {f"g_h{i}_relu": e.g_act[i] for i in range(4)}
        for i in range(4):
            acts[f"d_h{i}_conv"] = e.d_act[i][:B]
            acts[f"d_h{i}_conv_1"] = e.d_act[i][B:]
        for name, t in acts.items():
            s.histogram(name + "/activations", t.float().cpu().numpy())
            s.scalar(name + "/sparsity", float(K.zero_fraction(t.contiguous())))
        for name, t in e.params.items():
            s.histogram(name, t.detach().cpu().numpy())
        return s.serialize()

    def state_dict(self):
        return self.engine.state_dict()

    def load_state_dict(self, sd):
        self.engine.load_state_dict(sd)
