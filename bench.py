"""Headline benchmark: DCGAN train images/sec (D+G step, device-timed, max over ranks).

Contract (driver): ``python bench.py --gpus N --steps K --warmup W`` (N>1: launched under
``torch.distributed.run``, one rank per GPU).  Prints ONE JSON line on rank 0.

Config = BASELINE.json config 2: DCGAN 64x64 (ngf=ndf=64, z=100), sync data-parallel, batch 64 per
GPU, bf16 compute, synthetic images, random-init weights.

Arms:
  --impl ours       the B200 engine (hand-written sm_100a kernels, CUDA-graph replay)   [default]
  --impl ddp        the PyTorch DDP + NCCL + cuDNN baseline of BASELINE.md (baseline/ddp_dcgan.py)
  --impl reference  the unmodified reference from baseline/_ref (TensorFlow 0.x / Python 2: cannot be
                    installed offline here -> prints {"impl": "reference", "unavailable": ...})

Timing: W >= 3 warm-up steps; every timed step has its own CUDA-event pair on the launching stream
and the L2 is flushed (256 MiB write) between steps, outside the pairs; the reported time is the sum
over the K steps, max over ranks.  The e2e figure goes through the public ``submit`` / ``result`` API
(``train_step`` split in two) with a pinned-host -> device copy of the step's inputs and a device -> host
read of the losses every step; the copy of step i+1 overlaps the compute of step i.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "DCGAN train images/sec (D+G step, device-timed max over ranks)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "ddp", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="images per GPU (reference: 64, hard-coded)")
    ap.add_argument("--comm", default="fused", choices=["fused", "nccl"], help="multi-GPU gradient path")
    ap.add_argument("--no-graph", action="store_true",
                    help="ours: launch the kernels eagerly.  ddp: eager PyTorch + the DDP wrapper (the round-1 bar)")
    ap.add_argument("--wide", action="store_true", help="128x128, ngf=ndf=128 (BASELINE.json config 3)")
    ap.add_argument("--sync-bn", action="store_true")
    ap.add_argument("--mode", default="sync", choices=["sync", "async"],
                    help="async = parameter-server mode over peer memory (BASELINE.json config 4)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap,index")

    def __init__(self, index):
        """``index``: one GPU index or a comma-separated list (every GPU of the job, sampled by rank 0)."""
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i",
                 str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.mark_lo, self.mark_hi = 0, None
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def mark_begin(self):
        """Start of the timed region: samples before this line count as idle."""
        self.mark_lo = max(len(self.lines) - max(str(self.index).count(",") + 1, 1), 0)

    def mark_end(self):
        time.sleep(0.12)  # let the sample that covers the end of the region arrive
        self.mark_hi = len(self.lines)

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, by_gpu, pw = [], [], set(), {}, {}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        lines = self.lines[self.mark_lo:self.mark_hi] if self.mark_hi else self.lines
        for ln in lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            gpu = parts[7] if len(parts) > 7 else "0"
            by_gpu.setdefault(gpu, []).append(float(parts[0]))
            try:
                pw.setdefault(gpu, []).append(float(parts[2]))
            except ValueError:
                pass
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        out = {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
               "samples": len(sm), "reasons": sorted(reasons)}
        if len(by_gpu) > 1:  # the slowest GPU bounds a synchronous step: report every GPU's median under load
            out["sm_mhz"] = min(statistics.median(v) for v in by_gpu.values())
            out["sm_mhz_by_gpu"] = {k: statistics.median(v) for k, v in sorted(by_gpu.items(), key=lambda kv: int(kv[0]))}
            out["power_w_by_gpu"] = {k: round(statistics.median(v)) for k, v in sorted(pw.items(), key=lambda kv: int(kv[0]))}
        return out


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def reference_arm(args):
    """The reference is three Python-2 / TensorFlow-0.x scripts with no setup.py; see DESIGN.md."""
    why = None
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    try:
        sys.path.insert(0, ref_dir)
        import tensorflow  # noqa: F401
        why = "tensorflow imported but the reference needs the TF 0.x API (tf.train.Supervisor, tf.app.flags)"
    except Exception as e:  # pragma: no cover - depends on the image
        why = ("reference needs Python 2 + TensorFlow 0.7-0.8 (tf.train.Supervisor, tf.histogram_summary, "
               f"`print len(image)`); no setup.py/pyproject to pip-install and tensorflow is not installable "
               f"offline ({type(e).__name__})")
    rank, _, _ = dist_env()
    if rank == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


def multi_gpu_check(eng, comm, dist, torch, rank, world):
    """Driver-visible multi-GPU numerics, outside every timed region (N > 1, fused backend):
    (a) every rank holds bit-identical bf16 parameters after the timed steps;
    (b) one optimizer step through the fused reduce-scatter + Adam + all-gather kernel and one through
        NCCL all-reduce + the single-GPU Adam kernel, from identical state and rank-dependent gradients, agree."""
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as Kn
    A = eng.arena
    torch.cuda.synchronize()
    dist.barrier()
    # ---- (a) replicas in lock-step ------------------------------------------------------------------------
    lo, hi = getattr(comm, "pull_range", (0, 0))
    shadow = A.bf16.view(torch.int16).to(torch.int64)
    if hi > lo:   # gathered by its consumer at the next forward pass: only the owners' copies are current
        shadow = torch.cat([shadow[:lo], shadow[hi:]])
    mine = torch.stack([shadow.sum(), (shadow * (torch.arange(shadow.numel(), device=shadow.device) % 8191)).sum()])
    allv = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allv, mine)
    identical = all(bool((v == allv[0]).all()) for v in allv)
    # ---- (b) fused step vs NCCL + Adam --------------------------------------------------------------------
    keep = dict(master=A.master.clone(), m=A.m.clone(), v=A.v.clone(), bf16=A.bf16.clone(), cnt=eng.counters.clone())
    gen = torch.Generator(device=A.master.device).manual_seed(4321 + rank)
    grads = torch.randn(A.padded, device=A.master.device, generator=gen) * 1e-2
    eng.grads.copy_(grads)
    torch.cuda.synchronize(); dist.barrier()
    comm.optimizer_step(eng, "both")
    torch.cuda.synchronize(); dist.barrier()
    fused_master, fused_bf16 = A.master.clone(), A.bf16.clone()
    A.master.copy_(keep["master"]); A.m.copy_(keep["m"]); A.v.copy_(keep["v"]); A.bf16.copy_(keep["bf16"])
    ref_g = grads.clone()
    dist.all_reduce(ref_g)
    Kn.adam(A.master, ref_g, A.m, A.v, A.bf16, eng.counters, eng.lr, eng.beta1, eng.beta2, eng.eps, 1.0 / world)
    torch.cuda.synchronize()
    worst = 0.0
    for bucket in getattr(comm, "bucket_names", ("g", "d")):
        s0, s1 = comm.shard_range(bucket, comm.rank)      # the f32 master is current on its owner only
        if s1 > s0:
            d = (fused_master[s0:s1] - A.master[s0:s1]).abs().max().item()
            worst = max(worst, d / max(A.master[s0:s1].abs().max().item(), 1e-12))
    # bf16 shadows: compared on this rank's own shards (elsewhere this rank's f32 master / moments are stale by design
    # -- only the owner keeps them current -- so a locally recomputed shadow means nothing there); that the all-gather
    # then delivered the owners' values everywhere is check (c) below
    bf_diff = 0.0
    for bucket in getattr(comm, "bucket_names", ("g", "d")):
        s0, s1 = comm.shard_range(bucket, comm.rank)
        if s1 > s0:
            bf_diff = max(bf_diff, (fused_bf16[s0:s1].float() - A.bf16[s0:s1].float()).abs().max().item())
    # ---- (c) the fused step left bit-identical shadows on every rank ----------------------------------------
    sh2 = fused_bf16.view(torch.int16).to(torch.int64)
    if hi > lo:
        sh2 = torch.cat([sh2[:lo], sh2[hi:]])
    mine2 = torch.stack([sh2.sum(), (sh2 * (torch.arange(sh2.numel(), device=sh2.device) % 8191)).sum()])
    all2 = [torch.empty_like(mine2) for _ in range(world)]
    dist.all_gather(all2, mine2)
    identical = identical and all(bool((v == all2[0]).all()) for v in all2)
    # back to the state the benchmark left
    A.master.copy_(keep["master"]); A.m.copy_(keep["m"]); A.v.copy_(keep["v"]); A.bf16.copy_(keep["bf16"])
    eng.counters.copy_(keep["cnt"])
    t = torch.tensor([worst, bf_diff, 0.0 if identical else 1.0], device=A.master.device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    worst, bf_diff, bad = t.tolist()
    ok = bad == 0.0 and worst < 1e-5 and bf_diff < 1e-2
    return {"ranks_bit_identical": bad == 0.0, "fused_vs_nccl_adam_max_rel_err": worst,
            "bf16_param_max_abs_diff": bf_diff, "multicast": getattr(comm, "mc_grad", None) is not None, "ok": ok}


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist

    rank, world, local = dist_env()
    if args.gpus != world and rank == 0 and world > 1:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE {world}", file=sys.stderr)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    from distributed_tensorflow_for_dcgan_b200.config import ModelConfig
    B = args.batch
    if args.wide:
        cfg = ModelConfig(output_size=128, gf_dim=128, df_dim=128, batch_size=B)
    else:
        cfg = ModelConfig(batch_size=B)
    S = cfg.output_size
    K_steps, W = args.steps, max(args.warmup, 3)

    # synthetic data of the named shape, in pinned host memory (float32 NHWC like the reference feed)
    gen = torch.Generator().manual_seed(1234 + rank)
    n_host = 4
    host_real = [torch.empty(B, S, S, 3).uniform_(-1, 1, generator=gen).pin_memory() for _ in range(n_host)]
    host_z = [torch.empty(B, cfg.z_dim).uniform_(-1, 1, generator=gen).pin_memory() for _ in range(n_host)]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    launches_per_step = 0
    use_graph = False
    if args.impl == "ours":
        from distributed_tensorflow_for_dcgan_b200.engine import DCGANEngine
        from distributed_tensorflow_for_dcgan_b200.ops import _lib, kernels as Kn
        comm = None
        if world > 1:
            from distributed_tensorflow_for_dcgan_b200.parallel.sync_dp import make_comm
            if args.mode == "async":
                from distributed_tensorflow_for_dcgan_b200.parallel.async_ps import AsyncPSComm
                comm = AsyncPSComm(dev)
            else:
                comm = make_comm(args.comm, dev, sync_bn=args.sync_bn)
        eng = DCGANEngine(cfg, device=dev, seed=0, comm=comm)
        if comm is not None:
            comm.attach(eng)
        c0 = _lib.launch_count
        eng.real_f32.copy_(host_real[0]); eng.z.copy_(host_z[0])
        eng.step_body()
        launches_per_step = _lib.launch_count - c0
        torch.cuda.synchronize()
        use_graph = not args.no_graph and not (world > 1 and args.comm == "nccl")
        if use_graph:  # (NCCL collectives are launched eagerly: the baseline path is not graph-captured)
            eng.capture()

        def device_step(i):
            eng.run_step()

        def e2e_loop(n):
            # the public pipelined API: every step copies its own inputs from pinned host memory and has its
            # losses read back on the host; step i's read happens after step i+1 has been submitted
            prev, out = None, None
            for i in range(n):
                t = eng.submit(host_real[i % n_host], host_z[i % n_host])
                if prev is not None:
                    out = eng.result(prev)
                prev = t
            return eng.result(prev)

        def stage_inputs(i):
            eng.real_f32.copy_(host_real[i % n_host], non_blocking=True)
            eng.z.copy_(host_z[i % n_host], non_blocking=True)
        parallelism = f"dp{world}" + ("" if world == 1 else (f"-{args.comm}" if args.mode == "sync" else "-async-ps"))
        if args.sync_bn:
            parallelism += "-syncbn"
    else:
        from baseline.ddp_dcgan import BaselineTrainer
        use_graph = not args.no_graph
        if use_graph and world > 1:
            os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "0")  # NCCL collectives inside a capture
        tr = BaselineTrainer(dev, world=world, batch=B, s=S, gf=cfg.gf_dim, df=cfg.df_dim, z_dim=cfg.z_dim,
                             sync_bn=args.sync_bn, graph=use_graph)
        d_real = torch.empty(B, 3, S, S, device=dev).contiguous(memory_format=torch.channels_last)
        d_z = torch.empty(B, cfg.z_dim, device=dev)
        if use_graph:
            # the whole step (both backward passes, the all-reduces, both Adams) as one CUDA graph; every rank
            # must agree on whether the capture worked, or the collectives would no longer pair up
            d_real.uniform_(-1, 1); d_z.uniform_(-1, 1)
            ok = 1
            try:
                tr.capture(d_real, d_z)
            except Exception as e:   # pragma: no cover - depends on the torch / NCCL build
                print(f"rank {rank}: baseline graph capture failed ({type(e).__name__}: {e})", file=sys.stderr)
                ok = 0
            if world > 1:
                flag = torch.tensor([ok], device=dev)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                ok = int(flag.item())
            if not ok:
                raise SystemExit("baseline graph capture failed; rerun with --no-graph for the eager DDP bar")
        stage_nhwc = torch.empty(B, S, S, 3, device=dev)
        last = {}

        def stage_inputs(i):
            stage_nhwc.copy_(host_real[i % n_host], non_blocking=True)
            d_real.copy_(stage_nhwc.permute(0, 3, 1, 2))
            d_z.copy_(host_z[i % n_host], non_blocking=True)

        def device_step(i):
            last["l"] = tr.step(d_real, d_z)

        def e2e_loop(n):
            out = None
            for i in range(n):
                stage_inputs(i)
                dl, gl = tr.step(d_real, d_z)
                out = float(dl.item()), float(gl.item())
            return out
        parallelism = f"ddp{world}"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # clocks / throttle reasons are sampled from before the warm-up; only the samples taken during the two
    # timed regions are reported
    sampler = ClockSampler(",".join(str(i) for i in range(world)) if world > 1 else local)
    if rank == 0:
        sampler.start()
    # ---- warm-up ------------------------------------------------------------------------------------
    for i in range(W):
        stage_inputs(i)
        device_step(i)
    barrier()

    # ---- device-timed region: K steps, one event pair per step, L2 flushed in between -----------------
    if rank == 0:
        sampler.mark_begin()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K_steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(K_steps)]
    barrier()
    # Synchronous replicas end a step together (the exchange is a barrier), so a rank whose untimed inter-step work
    # (input copy, L2 flush) finishes EARLIER starts its next step earlier and then waits inside it: its event pair
    # would measure the other ranks' start skew, not work (profiles/README.md: ranks 4-7 read 6 % slower in every
    # synchronous arm while each of the eight GPUs alone runs the same 0.66 ms and the barrier-free async arm shows
    # no such pattern).  A one-element all-reduce in front of the start event lines the ranks up on the device; it is
    # outside the event pair and hides nothing: a genuinely slower rank still ends last and the max over ranks sees it.
    align = torch.zeros(1, device=dev) if (world > 1 and args.mode == "sync") else None
    for i in range(K_steps):
        stage_inputs(i)
        flush.fill_(i & 0xff)
        if align is not None:
            dist.all_reduce(align)
        starts[i].record()
        device_step(i)
        ends[i].record()
    barrier()
    per_step = [s.elapsed_time(e) for s, e in zip(starts, ends)]
    dev_ms = sum(per_step)

    # ---- end-to-end region through the public API: H2D inputs + D2H losses every step ----------------
    e2e_loop(3)
    barrier()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    out = e2e_loop(K_steps)
    t1.record()
    barrier()
    e2e_ms = t0.elapsed_time(t1)
    if rank == 0:
        sampler.mark_end()
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([dev_ms, e2e_ms], device=dev, dtype=torch.float64)
    per_rank = [round(dev_ms / K_steps, 4)]
    if world > 1:
        every = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(every, t)
        per_rank = [round(v[0].item() / K_steps, 4) for v in every]
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = t.tolist()
    mg_check = None
    if args.impl == "ours":
        from distributed_tensorflow_for_dcgan_b200.ops import kernels as Kn
        Kn.check_device_error()
        if world > 1 and args.mode == "sync" and args.comm == "fused":
            mg_check = multi_gpu_check(eng, comm, dist, torch, rank, world)
            Kn.check_device_error()

    if rank == 0:
        gb = B * world
        value = gb * K_steps / (dev_ms / 1e3)
        e2e_value = gb * K_steps / (e2e_ms / 1e3)
        base = None
        headline = not args.wide and args.mode == "sync" and B == 64 and not args.sync_bn
        try:
            with open(os.path.join(ROOT, "baseline", "measured.json")) as f:
                base = json.load(f).get("images_per_sec", {}).get(str(world))
        except Exception:
            pass
        h2d = B * S * S * 3 * 4 + B * cfg.z_dim * 4
        d2h = 3 * 4 + 2 * 8 if args.impl == "ours" else 8  # ours: 3 f32 losses + (adam_t, global_step) int64
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": "images/s", "n_gpus": world,
            "steps": K_steps, "warmup": W, "ms_per_step": round(dev_ms / K_steps, 4),
            "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (round(value / base, 3) if (base and headline) else None),
            "dtype": "bf16", "data": "synthetic (uniform [-1,1] images, random-init weights)",
            "impl": args.impl,
            "config": {"model": f"DCGAN {S}x{S} ngf={cfg.gf_dim} ndf={cfg.df_dim} z={cfg.z_dim}",
                       "global_batch": gb, "batch_per_gpu": B, "seq_len": None, "parallelism": parallelism,
                       "update_mode": "simultaneous", "replica_mode": args.mode, "cuda_graph": bool(use_graph),
                       "l2": "flushed between timed steps (256 MiB write outside the event pairs)",
                       "rank_alignment": ("one-element all-reduce before each start event (outside the event pair)"
                                          if (world > 1 and args.mode == "sync") else None),
                       "step_ms_min_med_max": [round(min(per_step), 4), round(statistics.median(per_step), 4),
                                               round(max(per_step), 4)],
                       "ms_per_step_by_rank": per_rank},
            "clocks": clocks,
            "e2e": {"value": round(e2e_value, 1), "unit": "images/s", "ms_per_step": round(e2e_ms / K_steps, 4),
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": launches_per_step * K_steps if args.impl == "ours" else 0,
            "launches_per_step": launches_per_step,
        }
        if mg_check is not None:
            line["multi_gpu_check"] = mg_check
        print(json.dumps(line))
    rc = 3 if (mg_check is not None and not mg_check["ok"]) else 0  # 3: a rank diverged or the fused exchange
    #                                     disagrees with NCCL + Adam -- the number does not count
    if world > 1:
        # Leave without tearing NCCL down: destroying a communicator that captured CUDA graphs still reference (the
        # graph-captured DDP arm) never returned on the 2-GPU box.  Everything is on the device and printed by now.
        sys.stdout.flush(); sys.stderr.flush()
        torch.cuda.synchronize()
        dist.barrier()
        os._exit(rc)
    return rc


if __name__ == "__main__":
    sys.exit(main())
