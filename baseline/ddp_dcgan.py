"""The number to beat: an equivalent PyTorch DDP + NCCL + cuDNN DCGAN.

Same architecture as the reference graph (``distriubted_model.py:83-128``: ngf=ndf=64, z=100,
64x64x3, 5x5 stride-2 layers, batch norm, 9,451,908 parameters), bf16 autocast, channels_last,
fused Adam, batch 64 per GPU, synthetic data, simultaneous D+G update from one forward pass
(``image_train.py:156-158``).  None of the framework's own kernels or engine is used here: this is
library code only (cuDNN / cuBLAS / NCCL through PyTorch).

Two ways to run the step:

* ``graph=True`` (the bar ``bench.py --impl ddp`` reports against): the whole step -- forward, both backward
  passes, the gradient all-reduces and both fused Adams -- is captured once into a CUDA graph and replayed, so
  the library kernels are not held back by Python / launch overhead.  With several GPUs the gradients live in
  one flat buffer per network and are averaged by two NCCL all-reduces inside the captured region (D's runs
  on NCCL's stream under G's backward pass), which is DDP's bucketed all-reduce written out by hand.
* ``graph=False``: eager PyTorch with the ``DistributedDataParallel`` wrapper (round-1 bar, kept for reference).
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn
import torch.nn.functional as F


class Generator(nn.Module):
    def __init__(self, z_dim=100, gf=64, s=64, c=3):
        super().__init__()
        self.s16, self.gf = s // 16, gf
        self.lin = nn.Linear(z_dim, gf * 8 * self.s16 * self.s16)
        self.bn0 = nn.BatchNorm2d(gf * 8, eps=1e-5, momentum=0.1)
        chans = [gf * 8, gf * 4, gf * 2, gf, c]
        self.deconvs = nn.ModuleList(
            nn.ConvTranspose2d(chans[i], chans[i + 1], 5, 2, 2, output_padding=1) for i in range(4))
        self.bns = nn.ModuleList(nn.BatchNorm2d(chans[i + 1], eps=1e-5, momentum=0.1) for i in range(3))

    def forward(self, z):
        h = self.lin(z).view(-1, self.gf * 8, self.s16, self.s16).contiguous(memory_format=torch.channels_last)
        h = F.relu(self.bn0(h))
        for i in range(3):
            h = F.relu(self.bns[i](self.deconvs[i](h)))
        return torch.tanh(self.deconvs[3](h))


class Discriminator(nn.Module):
    def __init__(self, df=64, s=64, c=3):
        super().__init__()
        chans = [c, df, df * 2, df * 4, df * 8]
        self.convs = nn.ModuleList(nn.Conv2d(chans[i], chans[i + 1], 5, 2, 2) for i in range(4))
        self.bns = nn.ModuleList(nn.BatchNorm2d(chans[i + 1], eps=1e-5, momentum=0.1) for i in range(1, 4))
        self.lin = nn.Linear(df * 8 * (s // 16) ** 2, 1)

    def forward(self, x):
        h = F.leaky_relu(self.convs[0](x), 0.2)
        for i in range(1, 4):
            h = F.leaky_relu(self.bns[i - 1](self.convs[i](h)), 0.2)
        return self.lin(h.flatten(1))


class DPair(nn.Module):
    """D on the real batch and on the fake batch in one wrapped forward (separate BN statistics, as in
    the reference's two ``discriminator`` calls)."""

    def __init__(self, d):
        super().__init__()
        self.d = d

    def forward(self, real, fake):
        return self.d(real), self.d(fake)


class BaselineTrainer:
    def __init__(self, device, world=1, batch=64, s=64, gf=64, df=64, z_dim=100, lr=2e-4, beta1=0.5,
                 sync_bn=False, graph=False):
        self.device, self.world, self.batch, self.z_dim, self.s = device, world, batch, z_dim, s
        self.graph_mode = graph
        self.manual_allreduce = graph and world > 1   # hand-written bucketed all-reduce instead of the DDP wrapper
        g = Generator(z_dim, gf, s).to(device).to(memory_format=torch.channels_last)
        d = DPair(Discriminator(df, s)).to(device).to(memory_format=torch.channels_last)
        if sync_bn and world > 1:
            g = nn.SyncBatchNorm.convert_sync_batchnorm(g)
            d = nn.SyncBatchNorm.convert_sync_batchnorm(d)
        if world > 1 and not self.manual_allreduce:
            from torch.nn.parallel import DistributedDataParallel as DDP
            g = DDP(g, device_ids=[device.index], gradient_as_bucket_view=True)
            d = DDP(d, device_ids=[device.index], gradient_as_bucket_view=True)
        self.g, self.d = g, d
        self.g_params = [p for p in g.parameters()]
        self.d_params = [p for p in d.parameters()]
        if self.manual_allreduce:
            import torch.distributed as dist
            for p in self.g_params + self.d_params:   # DDP's initial broadcast
                dist.broadcast(p.data, src=0)
        self.flat_g = self.flat_d = None
        if graph:
            self.flat_g = self._flat_grads(self.g_params)
            self.flat_d = self._flat_grads(self.d_params)
        self.opt_g = torch.optim.Adam(self.g_params, lr=lr, betas=(beta1, 0.999), fused=True, capturable=graph)
        self.opt_d = torch.optim.Adam(self.d_params, lr=lr, betas=(beta1, 0.999), fused=True, capturable=graph)
        self._graph = None

    @staticmethod
    def _flat_grads(params):
        """One flat gradient buffer per network; every ``p.grad`` is a view with the parameter's own strides
        (channels_last filters stay channels_last), so one all-reduce covers the network."""
        flat = torch.zeros(sum(p.numel() for p in params), device=params[0].device, dtype=params[0].dtype)
        off = 0
        for p in params:
            n = p.numel()
            p.grad = flat[off:off + n].as_strided(p.size(), p.stride())
            off += n
        return flat

    def n_params(self):
        return sum(p.numel() for p in self.g_params) + sum(p.numel() for p in self.d_params)

    def capture(self, real, z):
        """Warm up and capture the step on ``real`` / ``z`` (static input buffers); later ``step`` calls replay."""
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(3):
                self._step(real, z)
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._static_out = self._step(real, z)
        self._static_in = (real, z)
        return self._graph

    def step(self, real, z):
        if self._graph is not None:
            assert real is self._static_in[0] and z is self._static_in[1], "the captured step reads its static inputs"
            self._graph.replay()
            return self._static_out
        return self._step(real, z)

    def _step(self, real, z):
        """real: [B,3,s,s] channels_last bf16/f32 on device, z: [B,z_dim]."""
        if self.flat_g is not None:
            self.flat_g.zero_()
            self.flat_d.zero_()
        else:
            self.opt_g.zero_grad(set_to_none=True)
            self.opt_d.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            fake = self.g(z)
            lr_, lf_ = self.d(real, fake)
            ones, zeros = torch.ones_like(lr_), torch.zeros_like(lf_)
            d_loss = F.binary_cross_entropy_with_logits(lr_.float(), ones) + \
                F.binary_cross_entropy_with_logits(lf_.float(), zeros)
            g_loss = F.binary_cross_entropy_with_logits(lf_.float(), ones)
        # simultaneous update from one forward: d_loss -> D variables, g_loss -> G variables
        d_loss.backward(retain_graph=True, inputs=self.d_params)
        work_d = None
        if self.manual_allreduce:   # D's gradients are complete: average them under G's backward pass
            import torch.distributed as dist
            work_d = dist.all_reduce(self.flat_d, op=dist.ReduceOp.AVG, async_op=True)
        g_loss.backward(inputs=self.g_params)
        if self.manual_allreduce:
            dist.all_reduce(self.flat_g, op=dist.ReduceOp.AVG)
            work_d.wait()
        self.opt_d.step()
        self.opt_g.step()
        return d_loss, g_loss
