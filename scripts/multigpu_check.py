"""Multi-GPU correctness + bandwidth checks for the fused communication kernels.
Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 scripts/multigpu_check.py [--sweep]"""
import os, sys, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from distributed_tensorflow_for_dcgan_b200.config import ModelConfig
from distributed_tensorflow_for_dcgan_b200.engine import DCGANEngine
from distributed_tensorflow_for_dcgan_b200.ops import _lib, kernels as K
from distributed_tensorflow_for_dcgan_b200.parallel.sync_dp import NcclComm
from distributed_tensorflow_for_dcgan_b200.parallel.fused_comm import FusedComm
from distributed_tensorflow_for_dcgan_b200.parallel.async_ps import AsyncPSComm

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
results = {"world": world}


def log(*a):
    if rank == 0:
        print(*a, flush=True)


def make_inputs(B, seed):
    g = torch.Generator().manual_seed(seed)
    real = torch.empty(B, 64, 64, 3).uniform_(-1, 1, generator=g).to(dev)
    z = torch.empty(B, 100).uniform_(-1, 1, generator=g).to(dev)
    return real, z


def run_steps(eng, steps, B):
    out = None
    for s in range(steps):
        real, z = make_inputs(B, 1000 * s + rank)
        eng.real_f32.copy_(real); eng.z.copy_(z)
        eng.step_body()
    torch.cuda.synchronize()
    return eng.losses.tolist()


B = 16
cfg = ModelConfig(batch_size=B)
# ---- 1. fused RS+Adam+AG == NCCL all-reduce + Adam ----------------------------------------------------------
for mc in (False, True):
    ref = DCGANEngine(cfg, device=dev, seed=0, comm=(c0 := NcclComm(dev)))
    c0.attach(ref)
    fc = FusedComm(dev, use_multicast=mc)
    fus = DCGANEngine(cfg, device=dev, seed=0, comm=fc)
    fc.attach(fus)
    has_mc = fc.mc_grad is not None
    # identical state, rank-dependent gradients -> one optimizer step through each backend.  (Comparing
    # training trajectories instead would be ill-conditioned: Adam's first steps are sign-like, so
    # summation-order noise in near-zero gradients flips updates by +-lr.)
    A = fus.arena
    for step in range(1, 3):
        gen = torch.Generator(device=dev).manual_seed(100 * step + rank)
        gr = torch.randn(A.padded, device=dev, generator=gen) * 1e-2
        for e in (ref, fus):
            e.grads.copy_(gr)
            e.counters[0] = step
        torch.cuda.synchronize(); dist.barrier()
        ref.optimizer_step("both")
        fus.optimizer_step("both")
        torch.cuda.synchronize(); dist.barrier()
    K.check_device_error()
    dist.barrier()
    fus_state = fus.state_dict()   # reads the owners' shards of master / slots through peer memory
    ref_state = ref.state_dict()
    dist.barrier()
    worst = 0.0
    for k in ("g_h1/w", "d_h2_conv/w", "g_bn2/gamma", "d_h3_lin/Matrix", "g_h0_lin/Matrix", "d_h1_conv/w/Adam",
              "g_h3/w/Adam_1"):
        a, b = fus_state[k].float(), ref_state[k].float()
        err = (a - b).abs().max().item() / max(b.abs().max().item(), 1e-12)
        worst = max(worst, err)
    # parameters that kernels read in f32 must be current on EVERY rank (not only on their owner); and the
    # consumer-side gather of g_h0_lin/Matrix must have delivered the owners' values into the local shadow
    f32_ok = True
    for key, end in (("g", A.g_end), ("d", A.padded)):
        a0 = A.f32_start[key]
        ok_k = bool(torch.allclose(fus.arena.master[a0:end], ref.arena.master[a0:end], rtol=1e-4, atol=2e-6))  # a stale value would be off by ~lr = 2e-4; summation order alone gives ~1e-7
        if not ok_k:
            d = (fus.arena.master[a0:end] - ref.arena.master[a0:end]).abs()
            j = int(d.argmax())
            print(f"[rank {rank}] f32 mismatch bucket {key}: max abs diff {d.max().item():.3e} at +{j} "
                  f"(fused {fus.arena.master[a0 + j].item():.6e} ref {ref.arena.master[a0 + j].item():.6e}) "
                  f"n_bad {(d > 1e-6).sum().item()} of {end - a0}", flush=True)
        f32_ok &= ok_k
    fus.z.uniform_(-1, 1)
    fus.generator_forward(train=True)
    torch.cuda.synchronize(); dist.barrier()
    bf_diff = (fus.arena.bf16[:A.numel].float() - ref.arena.bf16[:A.numel].float()).abs().max().item()
    # every rank must hold identical bf16 parameters
    mine = fus.arena.bf16.float().sum().item()
    allv = [None] * world
    dist.all_gather_object(allv, mine)
    log(f"[fused vs nccl] multicast_requested={mc} multicast_used={has_mc} worst_rel_err={worst:.3e} "
        f"bf16_maxdiff={bf_diff:.3e} ranks_identical={len(set(allv)) == 1}")
    results[f"fused_mc{int(mc)}"] = dict(multicast_used=has_mc, worst_rel_err=worst, bf16_maxdiff=bf_diff,
                                         ranks_identical=len(set(allv)) == 1)
    assert f32_ok, "f32-consumed parameters are stale on a non-owner rank"
    assert worst < 1e-5 and bf_diff < 1e-2 and len(set(allv)) == 1, "fused comm mismatch"
    # and a few real training steps stay finite and in lock-step
    l_fus = run_steps(fus, 3, B)
    K.check_device_error()
    log("  training losses", l_fus)
    # graph capture + replay with the side-stream bucket
    fus.capture()
    for s in range(3):
        real, z = make_inputs(B, 77 + s + rank)
        fus.train_step(real, z)
    K.check_device_error()
    log("  graph replay ok, step", int(fus.counters[1].item()))
    del ref, fus, fc, c0
    torch.cuda.synchronize(); dist.barrier()

# ---- 2. cross-GPU batch-norm statistics ------------------------------------------------------------------------
fc = FusedComm(dev)
buf = torch.randn(3, 2, 512, device=dev) + rank
expect = buf.clone(); dist.all_reduce(expect); expect /= world
for _ in range(3):
    b2 = buf.clone()
    fc.allreduce_bn(b2)
torch.cuda.synchronize()
err = (b2 - expect).abs().max().item()
log(f"[allreduce_bn] max err {err:.3e}")
assert err < 1e-5
results["allreduce_bn_err"] = err
sbn = DCGANEngine(cfg, device=dev, seed=0, comm=(fcs := FusedComm(dev, sync_bn=True)))
fcs.attach(sbn)
l = run_steps(sbn, 2, B)
K.check_device_error()
log("[sync_bn engine] losses", l)
del sbn, fcs, fc
torch.cuda.synchronize(); dist.barrier()

# ---- 3. async parameter-server mode -------------------------------------------------------------------------------
ac = AsyncPSComm(dev)
aeng = DCGANEngine(cfg, device=dev, seed=0, comm=ac)
ac.attach(aeng)
steps = 4 + rank  # deliberately different numbers of pushes per worker: nobody waits for anybody
run_steps(aeng, steps, B)
K.check_device_error()
dist.barrier()
total = torch.tensor([steps], device=dev); dist.all_reduce(total)
gs = ac.global_step()
log(f"[async ps] pushes total={int(total)} global_step(rank0)={gs} shard_t={int(ac.shard_counters[0])}")
if rank == 0:
    assert gs == int(total), "lost global_step increments"
assert int(ac.shard_counters[0]) == int(total)
if rank == 0:
    ac.gather_state(aeng)   # one-sided: only the chief takes part
    assert torch.isfinite(aeng.arena.master).all()
dist.barrier()
results["async_total_pushes"] = int(total)
del aeng, ac
torch.cuda.synchronize(); dist.barrier()

# ---- 4. bandwidth sweep: fused RS+Adam+AG vs NCCL all-reduce + Adam kernel ------------------------------------------
if "--sweep" in sys.argv:
    import ctypes
    from distributed_tensorflow_for_dcgan_b200.parallel.fused_comm import _ptr_array
    sizes = [4800, 204800, 819200, 3276800, 9453568, 9453568 * 4]
    sweep = []
    fc = FusedComm(dev)
    for n in sizes:
        n = (n + 4 * world - 1) // (4 * world) * (4 * world)
        g = fc._symmetric(f"g{n}", n, torch.float32); g.normal_()
        bf = fc._symmetric(f"b{n}", n, torch.bfloat16)
        p = torch.randn(n, device=dev); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
        counters = torch.ones(2, dtype=torch.int64, device=dev)
        gp, bp = _ptr_array(fc.handles[f"g{n}"].buffer_ptrs), _ptr_array(fc.handles[f"b{n}"].buffer_ptrs)
        hg, hb = fc.handles[f"g{n}"], fc.handles[f"b{n}"]
        row = {"n": n}
        for name, mcg, mcb in (("p2p", None, None), ("nvls", int(hg.multicast_ptr) or None, int(hb.multicast_ptr) or None)):
            if name == "nvls" and not mcg:
                continue
            for blocks in (16, 32, 64, 148):
                def once():
                    _lib.call("dcgan_comm_rs_adam_ag", gp, bp, fc._sig_ptrs, None, p.data_ptr(), m.data_ptr(),
                              v.data_ptr(), counters.data_ptr(), fc.state[3].data_ptr(), rank, world, 0, n, 2e-4, 0.5,
                              0.999, 1e-8, 3, mcg, mcb, None, n, 0, 0, blocks)
                for _ in range(3): once()
                torch.cuda.synchronize(); dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(10): once()
                e1.record(); torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) / 10], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
                row[f"{name}_b{blocks}_us"] = round(t.item() * 1e3, 2)
        # NCCL baseline: all-reduce + local Adam over the full vector
        gl = torch.randn(n, device=dev); pb = torch.empty(n, dtype=torch.bfloat16, device=dev)
        def base():
            dist.all_reduce(gl)
            K.adam(p, gl, m, v, pb, counters, gscale=1.0 / world)
        for _ in range(3): base()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): base()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 10], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        row["nccl_allreduce_adam_us"] = round(t.item() * 1e3, 2)
        # bytes that must cross NVLink per rank: read (W-1)/W of n f32 grads, write (W-1)/W of n bf16 params
        row["nvlink_bytes_per_rank"] = int(n * (world - 1) / world * 6)
        sweep.append(row)
        log("[sweep]", json.dumps(row))
    results["sweep"] = sweep
    K.check_device_error()

if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/multigpu_check_w{world}.json", "w") as f:
        json.dump(results, f, indent=1)
    print("MULTIGPU CHECK PASSED")
dist.barrier()
dist.destroy_process_group()
