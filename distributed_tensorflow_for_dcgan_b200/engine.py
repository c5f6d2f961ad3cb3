"""B200 training engine: the whole D+G step as a fixed sequence of hand-written sm_100a kernels.

What the reference does in one ``sess.run([d_loss_fake, d_loss_real, g_loss, d_optim, g_optim,
global_step])`` (``image_train.py:156-158``) is restated here B200-first:

* every buffer (f32 master weights in TensorFlow layouts, bf16 compute copies, gradients, Adam
  slots, EMA shadows, activations) is allocated once, so the step is pointer-stable and can be
  replayed from a CUDA graph;
* forward and backward are written out by hand (no autograd on the hot path): D runs once on the
  concatenated [real; fake] batch with per-half batch-norm statistics, and its data-gradient pass
  carries three slabs -- d_loss through real, d_loss through fake, g_loss through fake -- so the
  "D(fake) is back-propagated twice" cost of the reference graph becomes one
  wider GEMM per layer;
* D and G are updated simultaneously from the same forward pass, as in the reference
  (``update_mode='simultaneous'``); ``'alternating'`` re-runs D(fake) after the D update.

The optimizer step is pluggable (``self.comm``): single GPU -> fused Adam kernel; multi GPU ->
the fused reduce-scatter + Adam + all-gather kernel over NVLink peer memory (parallel/sync_dp.py).
"""
from __future__ import annotations

import contextlib
import os
import types
from typing import Dict, List, Optional, Tuple

import torch

from .config import ModelConfig
from .models import dcgan, manifest
from .ops import kernels as K


class ParamArena:
    """Flat f32 master / grad / Adam buffers with TF-named, TF-layout views."""

    def __init__(self, cfg: ModelConfig, device, pad_to: int = 1024, alloc=None):
        self.specs = manifest.trainable_specs(cfg)
        self.offsets: Dict[str, Tuple[int, Tuple[int, ...]]] = {}
        # Arena order (independent of the checkpoint order): [G | D]; inside each, the big filters that the
        # tensor-core kernels read through the bf16 shadow come first, then everything the kernels read in
        # f32 (biases, batch-norm parameters, the 3-channel filters, d_h3_lin).  The multi-GPU optimizer
        # kernel all-gathers bf16 for the whole range and additionally f32 for the second part only.
        def group(pred):
            return [s for s in self.specs if pred(s.name)]
        g_all = group(manifest.is_g_var)
        d_all = group(manifest.is_d_var)
        off = 0
        self.f32_start = {}
        for key, specs in (("g", g_all), ("d", d_all)):
            if key == "d":
                self.g_end = off
            for part in (True, False):
                if not part:
                    self.f32_start[key] = off
                # g_h0_lin/Matrix goes last among the generator's big filters: its gradient is the last one of the
                # step, so [g_h1/w .. g_h3/w] can be updated early (engine: generator_backward) and the final
                # optimizer launch covers one contiguous range [g_h0_lin/Matrix, g_end)
                for s in sorted(specs, key=lambda t: t.name == "g_h0_lin/Matrix"):
                    if self.reads_bf16(s.name) == part:
                        self.offsets[s.name] = (off, s.shape)
                        off += (s.numel + 7) // 8 * 8  # every tensor 32-byte aligned (vector loads, TMA)
        self.numel = off
        self.padded = (off + pad_to - 1) // pad_to * pad_to
        if alloc is None:
            alloc = lambda name, numel, dtype: torch.zeros(numel, dtype=dtype, device=device)
        self.master = alloc("master", self.padded, torch.float32)
        self.m = alloc("adam_m", self.padded, torch.float32)
        self.v = alloc("adam_v", self.padded, torch.float32)
        self.bf16 = alloc("bf16", self.padded, torch.bfloat16)

    @staticmethod
    def reads_bf16(name: str) -> bool:
        """Variables consumed only through the bf16 shadow (the implicit-GEMM filters and g_h0_lin/Matrix)."""
        return name == "g_h0_lin/Matrix" or (name.endswith("/w") and name not in ("g_h4/w", "d_h0_conv/w"))

    def view(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        off, shape = self.offsets[name]
        n = 1
        for s in shape:
            n *= s
        return flat[off:off + n].view(shape)

    def named(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {s.name: self.view(flat, s.name) for s in self.specs}


class DCGANEngine:
    def __init__(self, cfg: ModelConfig = ModelConfig(), device="cuda", seed: int = 0, lr: float = 2e-4,
                 beta1: float = 0.5, beta2: float = 0.999, eps: float = 1e-8, update_mode: str = "simultaneous",
                 comm=None, init: Optional[Dict[str, torch.Tensor]] = None):
        assert update_mode in ("simultaneous", "alternating")
        from .ops import _lib
        _lib.load()  # fail loudly if the kernel library is missing
        self.cfg, self.device = cfg, torch.device(device)
        self.lr, self.beta1, self.beta2, self.eps = lr, beta1, beta2, eps
        self.update_mode = update_mode
        self.comm = comm
        self.B = B = cfg.batch_size
        # the communication backend may hand out peer-mapped (symmetric) memory for the buffers it touches
        alloc = None
        if comm is not None and hasattr(comm, "alloc"):
            alloc = lambda name, numel, dtype: comm.alloc(name, numel, dtype)
        pad_to = 1024 * (comm.world if comm is not None and hasattr(comm, "world") else 1)
        self.arena = A = ParamArena(cfg, self.device, pad_to=pad_to, alloc=alloc)
        self.params = A.named(A.master)
        src = init if init is not None else dcgan.init_params(cfg, seed=seed)
        for n, t in self.params.items():
            t.copy_(src[n].to(self.device, torch.float32))
        dev = self.device
        f32 = dict(dtype=torch.float32, device=dev)
        bf = dict(dtype=torch.bfloat16, device=dev)

        # ---- layer geometry ---------------------------------------------------------------------------
        s, gf, df = cfg.output_size, cfg.gf_dim, cfg.df_dim
        self.g_ch = [gf * 8, gf * 4, gf * 2, gf, cfg.c_dim]
        self.g_sz = [s // 16, s // 8, s // 4, s // 2, s]
        self.d_ch = [cfg.c_dim, df, df * 2, df * 4, df * 8]
        self.d_sz = [s, s // 2, s // 4, s // 8, s // 16]
        assert cfg.c_dim == 3, "the 3-channel kernels assume RGB"

        # ---- zero-per-step region: gradients + BN reduction scratch, one memset ----------------------
        def site_sizes(C, groups, slabs):
            return groups * 2 * C, slabs * 2 * C
        scratch = 0
        self._bn_off = {}
        for i in range(4):
            a, b = site_sizes(self.g_ch[i], 1, 1)
            self._bn_off[f"g_bn{i}"] = (scratch, scratch + a)
            scratch += a + b
        for i in (1, 2, 3):
            a, b = site_sizes(self.d_ch[i + 1], 2, 3)
            self._bn_off[f"d_bn{i}"] = (scratch, scratch + a)
            scratch += a + b
        self.zero_region = (alloc("zero_region", A.padded + scratch, torch.float32) if alloc is not None
                            else torch.zeros(A.padded + scratch, **f32))
        self.grads = self.zero_region[:A.padded]
        self.grad_views = A.named(self.grads)
        self._scratch = self.zero_region[A.padded:]

        def bn_bufs(name, C, groups, slabs):
            o1, o2 = self._bn_off[name]
            sums = self._scratch[o1:o1 + groups * 2 * C].view(groups, 2, C)
            red = self._scratch[o2:o2 + slabs * 2 * C].view(slabs, 2, C)
            return sums, red
        self.bn: Dict[str, dict] = {}
        self.ema_flat_parts: List[torch.Tensor] = []
        for i in range(4):
            C = self.g_ch[i]
            sums, red = bn_bufs(f"g_bn{i}", C, 1, 1)
            self.bn[f"g_bn{i}"] = dict(C=C, groups=1, slabs=1, sums=sums, red=red,
                                       saved=torch.zeros(1, 2, C, **f32), ema_mean=torch.zeros(1, C, **f32),
                                       ema_var=torch.zeros(1, C, **f32), act="relu")
        for i in (1, 2, 3):
            C = self.d_ch[i + 1]
            sums, red = bn_bufs(f"d_bn{i}", C, 2, 3)
            self.bn[f"d_bn{i}"] = dict(C=C, groups=2, slabs=3, sums=sums, red=red,
                                       saved=torch.zeros(2, 2, C, **f32), ema_mean=torch.zeros(2, C, **f32),
                                       ema_var=torch.zeros(2, C, **f32), act="lrelu")

        # ---- bf16 compute copies of the weights -------------------------------------------------------
        # one flat bf16 shadow in the TF layouts serves every use: K-major B for the transposed-conv mode,
        # MN-major B for the strided mode -- no per-use transposed copies
        self.w_native = A.named(A.bf16)

        # ---- activations ------------------------------------------------------------------------------
        self.z = torch.zeros(B, cfg.z_dim, **f32)
        self.real_f32 = torch.zeros(B, s, s, 3, **f32)       # landing buffer of the input pipeline (H2D)
        self.g_pre = [torch.zeros(B, self.g_sz[i], self.g_sz[i], self.g_ch[i], **bf) for i in range(4)]
        self.g_act = [torch.zeros_like(t) for t in self.g_pre]
        self.d_in = torch.zeros(2 * B, s, s, 3, **bf)        # [real; G(z)]
        self.d_act0 = torch.zeros(2 * B, self.d_sz[1], self.d_sz[1], self.d_ch[1], **bf)
        self.d_pre = [None] + [torch.zeros(2 * B, self.d_sz[i + 1], self.d_sz[i + 1], self.d_ch[i + 1], **bf)
                               for i in (1, 2, 3)]
        self.d_act = [self.d_act0] + [torch.zeros_like(t) for t in self.d_pre[1:]]
        self.logits = torch.zeros(2 * B, **f32)
        self.probs = torch.zeros(2 * B, **f32)
        self.losses = torch.zeros(3, **f32)
        self._losses_pass2 = torch.zeros(3, **f32)
        self.dlogit = torch.zeros(3 * B, **f32)
        self.counters = (alloc("counters", 2, torch.int64) if alloc is not None
                         else torch.zeros(2, dtype=torch.int64, device=dev))  # [adam_t, global_step]
        # gradients w.r.t. activations (3 slabs through D, 1 through G)
        self.dd_act = [torch.zeros(3 * B, *t.shape[1:], **bf) for t in self.d_act]
        self.dd_pre = [torch.zeros(3 * B, *t.shape[1:], **bf) for t in self.d_act]
        self.dimg = torch.zeros(B, s, s, 3, **bf)             # d g_loss / d (pre-tanh generator output)
        self.dg_act = [torch.zeros_like(t) for t in self.g_act]
        self.dg_pre = [torch.zeros_like(t) for t in self.g_act]
        # sampler buffers (EMA batch norm, train=False)
        self.s_pre = [torch.zeros_like(t) for t in self.g_pre]
        self.s_act = [torch.zeros_like(t) for t in self.g_pre]
        self.s_img = torch.zeros(B, s, s, 3, **bf)

        # filter gradients are off the critical path (only the optimizer consumes them): they run on a side
        # stream next to the data-gradient chain, filling the SMs that the small grids leave idle
        self.wgrad_overlap = os.environ.get("DCGAN_WGRAD_OVERLAP", "1") != "0"
        self.fuse_head_loss = os.environ.get("DCGAN_FUSE_HEAD_LOSS", "1") != "0"
        self.fuse_bn_bwd = (os.environ.get("DCGAN_FUSE_BN_BWD", "1") != "0"
                            and (int(os.environ.get("DCGAN_IGEMM_TUNING", "1")) & 128) == 0)
        # measured: 0.876 -> 0.815 ms/step on one GPU, 0.958 -> 0.905 on two (profiles/README.md)
        # batch norm inside the producing conv's launch (grid-wide barrier + elementwise tail), and one-launch
        # batch-norm kernels for the layers whose producer is not an implicit-GEMM conv.  All of them are issued on
        # the main stream only: two kernels that wait on a grid barrier must not share the device.
        # DCGAN_FUSE_BN is a bit mask for A/B runs: 1 forward conv tails, 2 backward conv tails, 4 one-launch forward
        # batch norm (g_bn0), 8 one-launch backward batch norm (d_bn3, g_bn3); the first-generation kernels have none.
        # Default off: measured on B200 (profiles/README.md, "batch norm inside the conv launch") every variant is
        # slower than the separate kernels -- with programmatic dependent launch a kernel boundary costs ~1.5 us, the
        # software grid barrier ~3 us, and the tail runs at the conv kernel's low occupancy.
        self.fuse_bn = int(os.environ.get("DCGAN_FUSE_BN", "0"))
        if (int(os.environ.get("DCGAN_IGEMM_TUNING", "1")) & 128) != 0:
            self.fuse_bn &= ~3
        self._grid_barrier = torch.zeros(2, dtype=torch.int32, device=self.device)
        self.fuse_act_bwd = (os.environ.get("DCGAN_FUSE_ACT_BWD", "1") != "0"
                             and (int(os.environ.get("DCGAN_IGEMM_TUNING", "1")) & 128) == 0)
        self.early_adam = os.environ.get("DCGAN_EARLY_ADAM", "1") != "0"
        self.memset_nodes = os.environ.get("DCGAN_MEMSET_NODES", "1") != "0"  # cudaMemsetAsync instead of fill kernels
        self.prioritize_main = os.environ.get("DCGAN_MAIN_PRIORITY", "1") != "0"
        self.split_d_forward = self.wgrad_overlap and os.environ.get("DCGAN_SPLIT_D_FORWARD", "0") != "0"
        self.side = torch.cuda.Stream(self.device)
        self._side_used = False
        # split-K scratch of the implicit-GEMM kernels (arrival counters + partial tiles): the chain and the side
        # stream run concurrently, so each gets its own (ops/_lib.set_workspace is per host thread)
        ws_bytes = 64 * 1024 * 1024 + 65536
        self._ws_main = torch.zeros(ws_bytes, dtype=torch.uint8, device=dev)
        self._ws_side = torch.zeros(ws_bytes, dtype=torch.uint8, device=dev)
        # set by callers that run a forward pass on ONE rank only (the chief's sample-time evaluation): cross-replica
        # batch-norm statistics are a collective and must not be entered alone
        self._local_bn_only = False
        # 1.0 freezes the EMA shadows for a pass (second D pass of the alternating update)
        self._ema_decay_override: Optional[float] = None
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self.refresh_compute_weights()

    # ------------------------------------------------------------------------------------------------
    # weights
    # ------------------------------------------------------------------------------------------------
    def refresh_compute_weights(self, cast: bool = True) -> None:
        """bf16 shadow of the master weights (flat cast; the optimizer kernels keep it current)."""
        A = self.arena
        if cast:
            K.cast_bf16(A.master, A.bf16)

    # ------------------------------------------------------------------------------------------------
    # forward
    # ------------------------------------------------------------------------------------------------
    def _stats_fusable(self, name, pre) -> bool:
        """The conv epilogue can accumulate the moments when every 128-row tile lies in one statistics group."""
        st = self.bn[name]
        rows = pre.numel() // st["C"]
        return (rows // st["groups"]) % 128 == 0

    def _bn_fwd(self, name, pre, act_out, train=True, stats_done=False):
        st = self.bn[name]
        C, G = st["C"], st["groups"]
        x2, y2 = pre.view(-1, C), act_out.view(-1, C)
        p = self.params
        if train:
            decay = self.cfg.bn_momentum if self._ema_decay_override is None else self._ema_decay_override
            if not stats_done and self._bn_tail_ok(4):
                K.bn_forward_fused(x2, y2, st["sums"], p[f"{name}/gamma"], p[f"{name}/beta"], st["saved"],
                                   st["ema_mean"], st["ema_var"], self._grid_barrier, G, st["act"], self.cfg.bn_eps,
                                   decay)
                return
            if not stats_done:
                K.bn_stats(x2, st["sums"], G)
            if self._sync_bn_active():
                self.comm.allreduce_bn(st["sums"])
            K.bn_apply(x2, y2, st["sums"], p[f"{name}/gamma"], p[f"{name}/beta"], st["saved"], st["ema_mean"],
                       st["ema_var"], G, st["act"], self.cfg.bn_eps, decay)
        else:
            K.bn_apply(x2, y2, st["sums"], p[f"{name}/gamma"], p[f"{name}/beta"], None, st["ema_mean"],
                       st["ema_var"], 1, st["act"], self.cfg.bn_eps, self.cfg.bn_momentum, use_ema=True)

    def _bn_tail_ok(self, kind: int) -> bool:
        """Batch norm may be finished inside the launch that produces its input: needs purely local statistics (a
        cross-replica reduction has to happen between the two halves) and the main stream (see ``fuse_bn``)."""
        return ((self.fuse_bn & kind) != 0 and not self._sync_bn_active()
                and torch.cuda.current_stream(self.device) != self.side)

    def _conv_bn_fwd(self, name, x, w, bias, pre, act_out, groups, transposed) -> None:
        """conv + batch norm (training mode) + activation: one launch when possible, else conv(+moments) and apply."""
        st, p = self.bn[name], self.params
        fused_stats = self._stats_fusable(name, pre)
        if fused_stats and self._bn_tail_ok(1):
            decay = self.cfg.bn_momentum if self._ema_decay_override is None else self._ema_decay_override
            K.conv_bn_fwd(x, w, bias, pre, act_out, st["sums"], p[f"{name}/gamma"], p[f"{name}/beta"], st["saved"],
                          st["ema_mean"], st["ema_var"], groups, st["act"], transposed, self.cfg.bn_eps, decay)
            return
        if fused_stats:
            (K.conv_tconv_stats if transposed else K.conv_fprop_kn_stats)(x, w, bias, pre, st["sums"], groups)
        elif transposed:
            K.conv_tconv(x, w, bias, out=pre)
        else:
            K.conv_fprop_kn(x, w, bias, out=pre)
        self._bn_fwd(name, pre, act_out, True, stats_done=fused_stats)

    def _sync_bn_active(self) -> bool:
        return (self.comm is not None and getattr(self.comm, "sync_bn", False) and not self._local_bn_only)

    def generator_forward(self, train: bool = True) -> torch.Tensor:
        """z -> images; training mode writes into the fake half of ``d_in``."""
        p, B = self.params, self.B
        pre = self.g_pre if train else self.s_pre
        act = self.g_act if train else self.s_act
        if train and self.comm is not None and hasattr(self.comm, "lin_fwd_gather"):
            # the parameter all-gather of the first generator layer happens inside its consumer: the kernel
            # pulls each slice of g_h0_lin/Matrix from the rank that owns (and just updated) it
            self.comm.lin_fwd_gather(self, pre[0].view(B, -1))
        else:
            K.lin_fwd(self.z, self.w_native["g_h0_lin/Matrix"], p["g_h0_lin/bias"], pre[0].view(B, -1))
        self._bn_fwd("g_bn0", pre[0], act[0], train)
        for i in (1, 2, 3):
            w = self.w_native[f"g_h{i}/w"].view(25, self.g_ch[i], self.g_ch[i - 1])
            if train:  # batch-norm moments fall out of the conv epilogue, the apply pass is the launch's tail
                self._conv_bn_fwd(f"g_bn{i}", act[i - 1], w, p[f"g_h{i}/biases"], pre[i], act[i], 1, True)
            else:
                K.conv_tconv(act[i - 1], w, p[f"g_h{i}/biases"], out=pre[i])
                self._bn_fwd(f"g_bn{i}", pre[i], act[i], False)
        out = self.d_in[B:] if train else self.s_img
        K.c3_tconv(act[3], p["g_h4/w"], p["g_h4/biases"], act="tanh", out=out, w_bf16=self.w_native["g_h4/w"])
        return out

    def discriminator_forward(self) -> None:
        p = self.params
        K.c3_fprop(self.d_in, p["d_h0_conv/w"], p["d_h0_conv/biases"], act="lrelu", out=self.d_act[0])
        for i in (1, 2, 3):
            w = self.w_native[f"d_h{i}_conv/w"].view(25, self.d_ch[i], self.d_ch[i + 1])
            self._conv_bn_fwd(f"d_bn{i}", self.d_act[i - 1], w, p[f"d_h{i}_conv/biases"], self.d_pre[i],
                              self.d_act[i], 2, False)
        K.head_fwd(self.d_act[3].view(2 * self.B, -1), p["d_h3_lin/Matrix"], p["d_h3_lin/bias"], self.logits)

    def discriminator_forward_half(self, h: int) -> None:
        """D on one half of the [real; fake] batch (h = 0 real, 1 fake).  The two calls of the reference's
        discriminator (image_train.py:80-82) share weights but not batch statistics, so the real half does not
        depend on the generator at all: the step runs it on the side stream underneath the generator's forward
        pass, whose small layers leave most SMs idle."""
        p, B = self.params, self.B
        sl = slice(h * B, (h + 1) * B)
        K.c3_fprop(self.d_in[sl], p["d_h0_conv/w"], p["d_h0_conv/biases"], act="lrelu", out=self.d_act[0][sl])
        for i in (1, 2, 3):
            name = f"d_bn{i}"
            st = self.bn[name]
            C = st["C"]
            w = self.w_native[f"d_h{i}_conv/w"].view(25, self.d_ch[i], self.d_ch[i + 1])
            pre, act = self.d_pre[i][sl], self.d_act[i][sl]
            sums = st["sums"][h:h + 1]
            if (pre.numel() // C) % 128 == 0:
                K.conv_fprop_kn_stats(self.d_act[i - 1][sl], w, p[f"d_h{i}_conv/biases"], pre, sums, 1)
            else:
                K.conv_fprop_kn(self.d_act[i - 1][sl], w, p[f"d_h{i}_conv/biases"], out=pre)
                K.bn_stats(pre.view(-1, C), sums, 1)
            K.bn_apply(pre.view(-1, C), act.view(-1, C), sums, p[f"{name}/gamma"], p[f"{name}/beta"],
                       st["saved"][h:h + 1], st["ema_mean"][h:h + 1], st["ema_var"][h:h + 1], 1, st["act"],
                       self.cfg.bn_eps, self.cfg.bn_momentum)
        K.head_fwd(self.d_act[3][sl].view(B, -1), p["d_h3_lin/Matrix"], p["d_h3_lin/bias"], self.logits[sl])

    # ------------------------------------------------------------------------------------------------
    # backward
    # ------------------------------------------------------------------------------------------------
    def _bnred_fusable(self, name, da, tile_rows_per_image) -> bool:
        """The data-gradient conv that produces ``da`` can form this layer's backward reductions in its epilogue
        when every 128-row tile of that conv lies in one slab (``DCGAN_FUSE_BN_BWD=0`` turns the fusion off)."""
        if not self.fuse_bn_bwd:
            return False
        imgs_per_slab = da.shape[0] // self.bn[name]["slabs"]
        return (imgs_per_slab * tile_rows_per_image) % 128 == 0

    def _dgrad_bnred(self, name, x, w, da, pre, transposed, dpre=None) -> str:
        """dgrad conv writing ``da`` + the batch-norm backward reductions of layer ``name`` in one launch ("red"), or
        -- statistics local, main stream -- that layer's whole backward pass (``dpre``, dgamma, dbeta; "full")."""
        st = self.bn[name]
        p, g = self.params, self.grad_views
        if dpre is not None and self._bn_tail_ok(2):
            K.conv_dgrad_bnbwd(x, w, da, pre, st["saved"], p[f"{name}/gamma"], p[f"{name}/beta"], st["red"], dpre,
                               g[f"{name}/gamma"], g[f"{name}/beta"], st["slabs"], st["groups"], st["groups"],
                               st["act"], transposed)
            return "full"
        K.conv_dgrad_bnred(x, w, da, pre, st["saved"], p[f"{name}/gamma"], p[f"{name}/beta"], st["red"],
                           st["slabs"], st["groups"], st["act"], transposed)
        return "red"

    def _bn_bwd(self, name, da, pre, dpre, reduce_done=False):
        st = self.bn[name]
        C = st["C"]
        p, g = self.params, self.grad_views
        slabs, G = st["slabs"], st["groups"]
        if reduce_done == "full":   # the conv that produced ``da`` finished this layer's backward pass as well
            return
        if not reduce_done and self._bn_tail_ok(8):
            K.bn_backward_fused(da.view(-1, C), pre.view(-1, C), dpre.view(-1, C), st["saved"], p[f"{name}/gamma"],
                                p[f"{name}/beta"], st["red"], g[f"{name}/gamma"], g[f"{name}/beta"],
                                self._grid_barrier, slabs, G, G, st["act"])
            return
        if not reduce_done:
            K.bn_bwd_reduce(da.view(-1, C), pre.view(-1, C), st["saved"], p[f"{name}/gamma"], p[f"{name}/beta"],
                            st["red"], slabs, G, st["act"])
        if self._sync_bn_active():
            self.comm.allreduce_bn(st["red"])
        K.bn_bwd_apply(da.view(-1, C), pre.view(-1, C), dpre.view(-1, C), st["saved"], p[f"{name}/gamma"],
                       p[f"{name}/beta"], st["red"], g[f"{name}/gamma"], g[f"{name}/beta"], slabs, G, G, st["act"])

    @contextlib.contextmanager
    def _off_critical_path(self):
        """Runs the enclosed launches on the side stream, ordered after everything issued so far on the current
        stream; ``_join_side()`` orders the current stream after them again."""
        if not self.wgrad_overlap:
            yield
            return
        from .ops import _lib
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.side.wait_event(ev)
        _lib.set_workspace(self._ws_side)
        try:
            with torch.cuda.stream(self.side):
                yield
        finally:
            _lib.set_workspace(self._ws_main)
        self._side_used = True

    def side_event(self) -> Optional[torch.cuda.Event]:
        """Event marking the end of the side-stream work issued so far (None when nothing is pending)."""
        if not self._side_used:
            return None
        ev = torch.cuda.Event()
        ev.record(self.side)
        return ev

    def _join_side(self) -> None:
        ev = self.side_event()
        if ev is not None:
            torch.cuda.current_stream(self.device).wait_event(ev)
            self._side_used = False

    def discriminator_backward(self, fused_loss: bool = False) -> None:
        """Three slabs of data gradient ([d_loss|real, d_loss|fake, g_loss|fake]); filter gradients from
        the first two only (``var_list=d_vars`` for d_loss, image_train.py:109-110).  ``fused_loss``: the
        losses / logit gradients / step counters (``K.gan_loss``) are produced by the same launch."""
        p, g, B = self.params, self.grad_views, self.B
        if fused_loss:
            K.head_loss_bwd(self.d_act[3].view(2 * B, -1), p["d_h3_lin/Matrix"], self.logits, self.losses,
                            self.dlogit, self.probs, self.counters, self.dd_act[3].view(3 * B, -1),
                            g["d_h3_lin/Matrix"], g["d_h3_lin/bias"], B, 1.0, True)
        else:
            K.head_bwd(self.d_act[3].view(2 * B, -1), p["d_h3_lin/Matrix"], self.dlogit,
                       self.dd_act[3].view(3 * B, -1), g["d_h3_lin/Matrix"], g["d_h3_lin/bias"], B)
        red_done = False   # d_bn3's upstream gradient comes from the head kernel, not from a conv
        for i in (3, 2, 1):
            self._bn_bwd(f"d_bn{i}", self.dd_act[i], self.d_pre[i], self.dd_pre[i], reduce_done=red_done)
            ci, co = self.d_ch[i], self.d_ch[i + 1]
            with self._off_critical_path():
                K.conv_wgrad(self.d_act[i - 1], self.dd_pre[i][:2 * B], g[f"d_h{i}_conv/w"].view(25, ci, co))
            w = self.w_native[f"d_h{i}_conv/w"].view(25, ci, co)
            # the data gradient lands in dd_act[i-1]; for i > 1 that is the input of d_bn{i-1}'s backward pass,
            # whose two reductions the conv's epilogue forms on the fly
            red_done = i > 1 and self._bnred_fusable(f"d_bn{i - 1}", self.dd_act[i - 1],
                                                     self.d_sz[i + 1] * self.d_sz[i + 1])
            act_done = False
            if red_done:
                red_done = self._dgrad_bnred(f"d_bn{i - 1}", self.dd_pre[i], w, self.dd_act[i - 1],
                                             self.d_pre[i - 1], True, dpre=self.dd_pre[i - 1])
            elif i == 1 and self.fuse_act_bwd and (B * self.d_sz[2] * self.d_sz[2]) % 128 == 0:
                # d_h0 is bias + LeakyReLU without batch norm: its activation backward and bias gradient ride in the
                # epilogue of d_h1's data gradient (slabs 0, 1 = d_loss feed the bias gradient, slab 2 = g_loss)
                K.conv_tconv_actbwd(self.dd_pre[1], w, self.dd_pre[0], self.d_act[0], g["d_h0_conv/biases"], 3, 2, 2,
                                    "lrelu")
                act_done = True
            else:
                K.conv_tconv(self.dd_pre[i], w, None, out=self.dd_act[i - 1])
        C0 = self.d_ch[1]
        if not act_done:
            K.act_bwd(self.dd_act[0].view(-1, C0), self.d_act[0].view(-1, C0), self.dd_pre[0].view(-1, C0),
                      g["d_h0_conv/biases"], "lrelu", bias_rows=2 * B * self.d_sz[1] * self.d_sz[1])
        with self._off_critical_path():
            K.c3_wgrad(self.d_in, self.dd_pre[0][:2 * B], g["d_h0_conv/w"])
        # g_loss slab only: gradient w.r.t. the generator's pre-tanh output
        K.c3_tconv(self.dd_pre[0][2 * B:], p["d_h0_conv/w"], None, aux=self.d_in[B:], out=self.dimg,
                   colsum=g["g_h4/biases"],  # g_h4's bias gradient falls out of the same pass, in f32
                   w_bf16=self.w_native["d_h0_conv/w"])

    def generator_backward(self, early_adam: bool = False, early_comm: bool = False) -> Optional[int]:
        """``early_adam``: the Adam update of g_h1/w .. g_h3/w (85 % of the generator's parameters) is issued on the
        side stream as soon as the last of their gradients is done and nothing reads those filters any more, under
        the tail of the backward pass; returns the arena offset where the remaining (final) update starts."""
        p, g, B = self.params, self.grad_views, self.B
        adam_from = None
        with self._off_critical_path():
            K.c3_wgrad(self.dimg, self.g_act[3], g["g_h4/w"])
        K.c3_fprop(self.dimg, p["g_h4/w"], None, out=self.dg_act[3])
        red_done = False   # g_bn3's upstream gradient comes from the 3-channel kernel
        for i in (3, 2, 1):
            self._bn_bwd(f"g_bn{i}", self.dg_act[i], self.g_pre[i], self.dg_pre[i], reduce_done=red_done)
            co, ci = self.g_ch[i], self.g_ch[i - 1]
            with self._off_critical_path():
                K.conv_wgrad(self.dg_pre[i], self.g_act[i - 1], g[f"g_h{i}/w"].view(25, co, ci))
            w = self.w_native[f"g_h{i}/w"].view(25, co, ci)
            red_done = self._bnred_fusable(f"g_bn{i - 1}", self.dg_act[i - 1], self.g_sz[i - 1] * self.g_sz[i - 1])
            if red_done:
                red_done = self._dgrad_bnred(f"g_bn{i - 1}", self.dg_pre[i], w, self.dg_act[i - 1],
                                             self.g_pre[i - 1], False, dpre=self.dg_pre[i - 1])
            else:
                K.conv_fprop_kn(self.dg_pre[i], w, None, out=self.dg_act[i - 1])
        if early_comm:
            # multi-GPU: the exchange + Adam of g_h1/w .. g_h3/w goes to the communication stream now, ordered after
            # g_h1's data gradient (last reader of those filters) and after the filter gradients on the side stream
            self.comm.launch_bucket(self, "g1", side_stream=True, also_wait=self.side_event())
        if early_adam and self.wgrad_overlap:
            A = self.arena
            lo, hi = A.offsets["g_h1/w"][0], A.offsets["g_h0_lin/Matrix"][0]
            if 0 <= lo < hi:
                # ordered after g_h1's data gradient (the last reader of these filters) and, on the side stream,
                # after the three filter gradients
                with self._off_critical_path():
                    self._adam_range(lo, hi)
                adam_from = hi
        self._bn_bwd("g_bn0", self.dg_act[0], self.g_pre[0], self.dg_pre[0], reduce_done=red_done)
        K.lin_wgrad(self.z, self.dg_pre[0].view(B, -1), g["g_h0_lin/Matrix"], g["g_h0_lin/bias"])
        return adam_from

    # ------------------------------------------------------------------------------------------------
    # optimizer
    # ------------------------------------------------------------------------------------------------
    def _adam_range(self, lo: int, hi: int) -> None:
        A = self.arena
        K.adam(A.master[lo:hi], self.grads[lo:hi], A.m[lo:hi], A.v[lo:hi], A.bf16[lo:hi], self.counters, self.lr,
               self.beta1, self.beta2, self.eps, 1.0)

    def optimizer_step(self, which: str = "both") -> None:
        A = self.arena
        self._join_side()  # every filter gradient must have landed
        if self.comm is not None:
            self.comm.optimizer_step(self, which)
        else:
            lo, hi = {"both": (0, A.padded), "g": (0, A.g_end), "d": (A.g_end, A.padded)}[which]
            self._adam_range(lo, hi)
        self.refresh_compute_weights(cast=False)

    # ------------------------------------------------------------------------------------------------
    # the step
    # ------------------------------------------------------------------------------------------------
    def _range(self, name: str):
        """NVTX range around a phase of the step when profiling is on (``engine.nvtx = True``, --profile)."""
        if getattr(self, "nvtx", False):
            return torch.cuda.nvtx.range(name)
        import contextlib
        return contextlib.nullcontext()

    def step_body(self) -> None:
        """Everything of one training iteration that runs on the device, in stream order."""
        from .ops import _lib
        _lib.set_workspace(self._ws_main)
        try:
            self._step_body()
        finally:
            _lib.set_workspace(None)

    def _step_body(self) -> None:
        B = self.B
        # gradients are not touched before the backward pass: their memset runs beside the forward pass
        if self.memset_nodes:
            with self._off_critical_path():
                K.zero_(self.grads)
            K.zero_(self._scratch)
        else:
            with self._off_critical_path():
                self.grads.zero_()
            self._scratch.zero_()
        deferred_g = (self.update_mode == "simultaneous" and self.comm is not None
                      and getattr(self.comm, "overlap", False) and getattr(self.comm, "deferred_g", False))
        if deferred_g:  # async mode: the previous step's generator gradients are pushed under this step's passes
            self.comm.launch_deferred_g(self)
        K.f32_to_bf16(self.real_f32, self.d_in[:B])
        split = self.split_d_forward and not (self.comm is not None and getattr(self.comm, "sync_bn", False))
        if split:  # D(real) needs nothing from G: it runs beside the generator's forward pass
            with self._off_critical_path():
                self.discriminator_forward_half(0)
        with self._range("G_forward"):
            self.generator_forward(train=True)
        with self._range("D_forward"):
            if split:
                self.discriminator_forward_half(1)
                self._join_side()
            else:
                self.discriminator_forward()
            if not self.fuse_head_loss:
                K.gan_loss(self.logits, self.losses, self.dlogit, self.probs, self.counters, B, 1.0, True)
        self._join_side()
        with self._range("D_backward"):
            self.discriminator_backward(fused_loss=self.fuse_head_loss)
        if self.update_mode == "simultaneous":
            if self.comm is not None and getattr(self.comm, "overlap", False):
                # D's gradient exchange + Adam runs on the communication stream under G's backward pass
                self.comm.launch_bucket(self, "d", side_stream=True, also_wait=self.side_event())
                early = getattr(self.comm, "has_early_generator_bucket", lambda: False)()
                self.generator_backward(early_comm=early)
                self._join_side()
                if deferred_g:
                    self.comm.join()
                    self.comm.stash_g(self)
                else:
                    self.comm.launch_bucket(self, "g2" if early else "g", side_stream=False)
                    self.comm.join()
                self.refresh_compute_weights(cast=False)
            elif self.comm is None and self.wgrad_overlap:
                # D's Adam (its gradients are complete) runs beside G's backward pass
                A = self.arena
                with self._off_critical_path():
                    self._adam_range(A.g_end, A.padded)
                adam_from = self.generator_backward(early_adam=self.early_adam)
                self._join_side()
                self._adam_range(adam_from or 0, A.g_end)
            else:
                self.generator_backward()
                self.optimizer_step("both")
        else:
            # D first, then G against the updated D (carpedm20-style alternation)
            self.optimizer_step("d")
            # The second D pass starts from clean reduction buffers: the batch-norm sums / backward reductions and
            # the generator-side gradients that the first pass already touched all accumulate (+=).  The EMA shadows
            # are not advanced a second time (CPUTrainer._alternating_step does the same).
            self._scratch.zero_()
            self.grads[:self.arena.g_end].zero_()
            self._ema_decay_override = 1.0
            try:
                self.discriminator_forward()
            finally:
                self._ema_decay_override = None
            # d_loss_real / d_loss_fake reported for the step are those of the first pass (before D's update), g_loss
            # the one G is trained on (against the updated D) -- CPUTrainer._alternating_step returns the same
            K.gan_loss(self.logits, self._losses_pass2, self.dlogit, self.probs, None, B, 1.0, False)
            self.losses[2:3].copy_(self._losses_pass2[2:3])
            self.discriminator_backward()
            self.generator_backward()
            self.optimizer_step("g")

    def capture(self) -> None:
        """Warm up on a side stream and capture the step into a CUDA graph."""
        state = self.snapshot_state()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                self.step_body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.restore_state(state)
        self._graph = torch.cuda.CUDAGraph()
        # Kernel nodes inherit the priority of the stream they were captured on: capturing the chain on a
        # high-priority stream lets the block scheduler serve it before the side-stream filler kernels.
        hp = torch.cuda.Stream(device=self.device, priority=-1) if self.prioritize_main else None
        with torch.cuda.graph(self._graph, stream=hp):
            self.step_body()
        self.restore_state(state)
        getattr(self.comm, "reset_deferred", lambda: None)()   # gradients stashed by the warm-up steps are not pushed

    def run_step(self) -> None:
        """One iteration on the inputs currently in ``real_f32`` / ``z`` (no host sync)."""
        if self._graph is not None:
            self._graph.replay()
        else:
            self.step_body()

    # -- pipelined public API ---------------------------------------------------------------------------
    # The reference feeds its step from an asynchronous input queue (tf.train.shuffle_batch,
    # image_input.py:41-46) and fetches the losses with the same sess.run (image_train.py:156-158).  Here the
    # host->device copy of a step's inputs runs on a copy stream into one of two staging slots while the
    # previous step computes, and the fetches land in pinned host slots: ``submit`` never blocks the host,
    # ``result`` waits only for its own step.
    def _pipe(self):
        st = getattr(self, "_pipe_state", None)
        if st is None:
            st = types.SimpleNamespace()
            st.copy_stream = torch.cuda.Stream(device=self.real_f32.device)
            st.stage_real = [torch.empty_like(self.real_f32) for _ in range(2)]
            st.stage_z = [torch.empty_like(self.z) for _ in range(2)]
            st.host_loss = [torch.empty(self.losses.numel(), dtype=self.losses.dtype).pin_memory() for _ in range(2)]
            st.host_cnt = [torch.empty(self.counters.numel(), dtype=self.counters.dtype).pin_memory()
                           for _ in range(2)]
            st.host_err = [torch.zeros(1, dtype=torch.int32).pin_memory() for _ in range(2)]
            st.ev_in = [torch.cuda.Event() for _ in range(2)]
            st.ev_consumed = [torch.cuda.Event() for _ in range(2)]
            st.ev_out = [torch.cuda.Event() for _ in range(2)]
            st.n = 0
            self._pipe_state = st
        return st

    def submit(self, real: torch.Tensor, z: torch.Tensor) -> int:
        """Enqueue one training step on ``real`` [B,S,S,3] / ``z`` [B,z_dim] (pinned-host, pageable or device
        tensors) and return a ticket for :meth:`result`.  Nothing here waits for the GPU.  Host inputs must stay
        untouched until ``result(ticket)`` returns; at most two tickets may be outstanding."""
        st = self._pipe()
        slot = st.n & 1
        main = torch.cuda.current_stream()
        # Host tensors go through the staging slot on the copy stream.  Device tensors are copied on the current
        # stream, which is the one their producer has been ordered against (e.g. DevicePrefetcher.next()).
        staged = [(src, stage, dst) for src, stage, dst in ((real, st.stage_real[slot], self.real_f32),
                                                            (z, st.stage_z[slot], self.z)) if not src.is_cuda]
        if staged:
            with torch.cuda.stream(st.copy_stream):
                st.copy_stream.wait_event(st.ev_consumed[slot])  # the step that last used this slot has read it
                for src, stage, _ in staged:
                    stage.copy_(src, non_blocking=True)
                st.ev_in[slot].record(st.copy_stream)
            main.wait_event(st.ev_in[slot])
            for _, stage, dst in staged:
                dst.copy_(stage, non_blocking=True)
            st.ev_consumed[slot].record(main)
        if real.is_cuda:
            self.real_f32.copy_(real, non_blocking=True)
        if z.is_cuda:
            self.z.copy_(z, non_blocking=True)
        self.run_step()
        st.host_loss[slot].copy_(self.losses, non_blocking=True)
        st.host_cnt[slot].copy_(self.counters, non_blocking=True)
        # the sticky device error flag travels with the fetches: a timed-out peer wait makes a rank skip its
        # shard's update, which must stop the job instead of silently desynchronising the replicas
        from .ops import _lib
        _lib.load().dcgan_copy_error_flag(st.host_err[slot].data_ptr(), _lib.stream_handle())
        st.ev_out[slot].record(main)
        st.n += 1
        return st.n - 1

    def result(self, ticket: int) -> Tuple[float, float, float, int]:
        """``(d_loss_real, d_loss_fake, g_loss, global_step)`` of a submitted step -- the fetches of
        ``image_train.py:156-158``."""
        st = self._pipe()
        if not (st.n - 2 <= ticket < st.n):
            raise ValueError(f"ticket {ticket} is no longer held (last submitted: {st.n - 1})")
        slot = ticket & 1
        st.ev_out[slot].synchronize()
        err = int(st.host_err[slot][0])
        if err != 0:
            from .ops import _lib
            raise _lib.KernelError(f"device-side error flag = {err} after step {ticket} (1: mbarrier wait timed "
                                   "out, 2: peer signal wait timed out): replicas may have diverged, aborting")
        l = st.host_loss[slot]
        return float(l[0]), float(l[1]), float(l[2]), int(st.host_cnt[slot][1])

    def train_step(self, real: torch.Tensor, z: torch.Tensor) -> Tuple[float, float, float, int]:
        """Public single-step API: copies the inputs (pinned host or device tensors), runs the step and
        reads back ``(d_loss_real, d_loss_fake, g_loss, global_step)``.  ``submit`` / ``result`` are the same
        thing split in two so that the next step's input copy overlaps this step's compute."""
        return self.result(self.submit(real, z))

    # ------------------------------------------------------------------------------------------------
    # sampler / state
    # ------------------------------------------------------------------------------------------------
    def sample(self, z: torch.Tensor) -> torch.Tensor:
        """``sampler(z)`` (distriubted_model.py:131-153): generator with EMA batch-norm statistics."""
        self.z.copy_(z)
        return self.generator_forward(train=False)

    def ema_named(self) -> Dict[str, torch.Tensor]:
        out = {}
        for i in range(4):
            m, v = manifest.ema_names(f"g_bn{i}")
            out[m], out[v] = self.bn[f"g_bn{i}"]["ema_mean"][0], self.bn[f"g_bn{i}"]["ema_var"][0]
        for i in (1, 2, 3):
            for gidx, sfx in enumerate(("", "_1")):
                m, v = manifest.ema_names(f"d_bn{i}{sfx}")
                out[m], out[v] = self.bn[f"d_bn{i}"]["ema_mean"][gidx], self.bn[f"d_bn{i}"]["ema_var"][gidx]
        return out

    def snapshot_state(self):
        A = self.arena
        return dict(master=A.master.clone(), m=A.m.clone(), v=A.v.clone(), counters=self.counters.clone(),
                    ema={k: t.clone() for k, t in self.ema_named().items()})

    def restore_state(self, st) -> None:
        A = self.arena
        A.master.copy_(st["master"]); A.m.copy_(st["m"]); A.v.copy_(st["v"])
        self.counters.copy_(st["counters"])
        for k, t in self.ema_named().items():
            t.copy_(st["ema"][k])
        self.refresh_compute_weights()

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """TF-named checkpoint contents: weights, EMA shadows, global_step, and the Adam slots under
        TensorFlow's slot names.  With sharded optimizer state (multi-GPU) this is a
        collective: every rank must call it."""
        A = self.arena
        if self.comm is not None and hasattr(self.comm, "flush"):
            self.comm.flush(self)   # async mode: the generator gradients still waiting for their deferred push
        if self.comm is not None and hasattr(self.comm, "gather_state"):
            self.comm.gather_state(self)
        sd = {n: t.detach().clone() for n, t in self.params.items()}
        sd.update({n: t.detach().clone() for n, t in self.ema_named().items()})
        async_ps = self.comm is not None and hasattr(self.comm, "shard_counters")
        if async_ps:   # async parameter-server mode: the applies of ALL workers, counted on the shard owners
            sd[manifest.GLOBAL_STEP_NAME] = torch.tensor(self.comm.global_step(), dtype=torch.int32)
        else:
            sd[manifest.GLOBAL_STEP_NAME] = self.counters[1].to(torch.int32).clone()
        m, v = A.named(A.m), A.named(A.v)
        for n in self.params:
            a, b = manifest.adam_slot_names(n)
            sd[a], sd[b] = m[n].clone(), v[n].clone()
        t = int(self.comm.shard_counters[0].item()) if async_ps else int(self.counters[0].item())
        sd["beta1_power"] = torch.tensor(self.beta1 ** (t + 1))
        sd["beta2_power"] = torch.tensor(self.beta2 ** (t + 1))
        sd["adam_t"] = torch.tensor(t, dtype=torch.int64) if async_ps else self.counters[0].clone()
        return sd

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        A = self.arena
        for n, t in self.params.items():
            t.copy_(sd[n].to(self.device))
        for n, t in self.ema_named().items():
            if n in sd:
                t.copy_(sd[n].to(self.device))
        m, v = A.named(A.m), A.named(A.v)
        for n in self.params:
            a, b = manifest.adam_slot_names(n)
            if a in sd:
                m[n].copy_(sd[a].to(self.device)); v[n].copy_(sd[b].to(self.device))
        if manifest.GLOBAL_STEP_NAME in sd:
            self.counters[1] = int(sd[manifest.GLOBAL_STEP_NAME])
        if "adam_t" in sd:
            self.counters[0] = int(sd["adam_t"])
        self.refresh_compute_weights()
