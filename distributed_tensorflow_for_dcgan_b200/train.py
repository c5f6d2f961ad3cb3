"""Training driver: role dispatch, supervisor, input pipeline and the step loop.

The reference packs cluster bootstrap, graph build, supervision and the loop into one 144-line
function (``train()``, image_train.py:51-194); here they are separate modules and this file only
wires them, keeping the observable behaviour: flag dump, mkdirs (``main``, :222-230), the log line
``Epoch: [%2d] step: [%2d] time: %4.4f, d_loss: %.8f, g_loss: %.8f`` (:160-162), the time-gated
summaries (:163-178), the chief's 8x8 sample grid whenever ``step % 100 == 1`` (:179-192) and the
``step < 1200000`` global budget (:150).
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch

from . import flags as F
from .config import steps_per_epoch
from .parallel import cluster as cl
from .supervisor import Supervisor
from .utils import sample as S

SAMPLE_SIZE = 64  # image_train.py:43


def _make_reader(FLAGS, data_dir, batch_size, cfg, seed, raw_f64=False):
    from .data import loader
    if FLAGS.synthetic_data:
        return loader.SyntheticImages(batch_size, cfg.output_size, cfg.c_dim, seed)
    kw = {}
    if FLAGS.shuffle_min_after_dequeue >= 0:
        kw = dict(min_after_dequeue=FLAGS.shuffle_min_after_dequeue,
                  capacity=FLAGS.shuffle_min_after_dequeue + 3 * batch_size)
    return loader.distorted_inputs(data_dir, batch_size, cfg.output_size, cfg.c_dim, seed=seed,
                                   num_threads=FLAGS.num_reader_threads,
                                   raw_f64=raw_f64, **kw)


def run_ps(FLAGS, role, cfg) -> int:
    """``server.join()`` (image_train.py:62-63)."""
    use_cuda = _use_cuda(FLAGS)
    if use_cuda:
        # GPU mode: the parameter shards live in the workers' peer-mapped HBM; nothing to serve.  The ps task is a
        # member of the gloo world only; it helps create the workers' NCCL group (collective) and then joins.
        cl.init_process_group(role, "gloo")
        import torch.distributed as dist
        cl.worker_group(role, "nccl")
        dist.barrier()   # workers started
        dist.barrier()   # workers finished
        print(f"ps {role.task_index}: joined; the workers own the parameter shards", flush=True)
        cl.shutdown()
        return 0
    from .parallel.ps import PSServer
    cl.init_process_group(role, "gloo")
    workers = [role.cluster.worker_rank(i) for i in range(role.cluster.num_workers)]
    server = PSServer(cfg, role.task_index, role.cluster.num_ps, workers, FLAGS.sync_replicas, FLAGS.learning_rate,
                      FLAGS.beta1, worker_timeout=FLAGS.worker_timeout_secs)
    applies = server.serve()
    print(f"ps {role.task_index}: served {applies} applies, global_step {server.global_step}", flush=True)
    # the receiver thread is still blocked in a receive from any worker, so the process group is not torn down
    # collectively: leave directly (the workers destroy their side on their own)
    sys.stdout.flush()
    os._exit(0)


def _use_cuda(FLAGS) -> bool:
    if FLAGS.device == "cpu":
        return False
    if FLAGS.device == "cuda":
        return True
    return torch.cuda.is_available()


def run_worker(FLAGS, role, cfg) -> int:
    use_cuda = _use_cuda(FLAGS)
    is_chief = role.is_chief
    batch = cfg.batch_size
    spe = steps_per_epoch(batch, FLAGS.train_size)
    ps_client = None
    comm = None
    if use_cuda:
        local = int(os.environ.get("LOCAL_RANK", role.task_index % max(torch.cuda.device_count(), 1)))
        torch.cuda.set_device(local)
        device = torch.device("cuda", local)
        wgroup = None
        if role.distributed:
            import torch.distributed as dist
            if role.cluster.num_ps:
                # ps tasks are CPU-only joiners: the world (ps + workers) meets over gloo, the workers get an NCCL
                # sub-group of their own for the symmetric-memory rendezvous and the initial broadcast
                cl.init_process_group(role, "gloo")
                wgroup = cl.worker_group(role, "nccl")
                dist.barrier()   # pairs with run_ps: workers started
            else:
                cl.init_process_group(role, "nccl", device=device)
                wgroup = dist.group.WORLD
            from .parallel.sync_dp import make_comm
            if FLAGS.sync_replicas:
                comm = make_comm(FLAGS.comm_backend, device, sync_bn=FLAGS.sync_bn, group=wgroup)
            else:
                from .parallel.async_ps import AsyncPSComm
                comm = AsyncPSComm(device, group=wgroup)
            if hasattr(comm, "set_peer_timeout"):
                comm.set_peer_timeout(FLAGS.peer_timeout_secs)
        if FLAGS.dtype == "fp32":
            # reference-precision path: the oracle model in f32 on the GPU through PyTorch's library kernels
            # (single process only); the hand-written engine computes in bf16
            if role.distributed:
                raise SystemExit("--dtype=fp32 runs the single-process oracle path; use bf16 for multi-GPU")
            from .trainer import CPUTrainer
            trainer = CPUTrainer(cfg, FLAGS.learning_rate, FLAGS.beta1, FLAGS.seed, FLAGS.update_mode, device=device)
        else:
            from .trainer import EngineTrainer
            trainer = EngineTrainer(cfg, FLAGS.learning_rate, FLAGS.beta1, FLAGS.seed, FLAGS.update_mode, device,
                                    comm, FLAGS.cuda_graph)
            trainer.engine.nvtx = bool(FLAGS.profile)   # NVTX ranges per phase for nsys / ncu range filters
    else:
        from .trainer import CPUTrainer
        allreduce_world = 1
        if role.distributed:
            cl.init_process_group(role, "gloo")
        trainer = CPUTrainer(cfg, FLAGS.learning_rate, FLAGS.beta1, FLAGS.seed, FLAGS.update_mode)
        if role.distributed and role.cluster.num_ps:
            from .parallel.ps import PSClient
            ps_client = PSClient(cfg, role.cluster.num_ps, trainer.params)
            trainer.ps = ps_client
        elif role.distributed:
            trainer.allreduce_world = role.cluster.num_workers

    sv = Supervisor(is_chief, FLAGS.checkpoint_dir, FLAGS.save_model_secs, FLAGS.save_summaries_secs,
                    FLAGS.max_steps)
    # ---- init or restore on the chief, everyone else waits for it (image_train.py:141-146) -----------------
    restored = sv.init_or_restore(trainer)
    if role.distributed:
        import torch.distributed as dist
        if ps_client is not None:
            if is_chief:
                ps_client.set_params(trainer.global_step, trainer.opt_d.t)
                if restored:
                    ps_client.set_slots(*trainer.adam_slots())
            # workers wait for the chief's initialisation, then pull it
            worker_group = _worker_barrier(role)
            ps_client.pull()
            trainer.global_step = ps_client.global_step
        elif not use_cuda:
            for t in trainer.params.values():
                dist.broadcast(t, src=role.cluster.worker_rank(0))
        elif comm is not None:
            # GPU path: every worker adopts the chief's (fresh or restored) variables, slots and counters -- the
            # "wait for the chief to initialise" of managed_session (image_train.py:141)
            eng = trainer.engine
            for t in (eng.arena.master, eng.arena.m, eng.arena.v, eng.counters):
                dist.broadcast(t, src=role.cluster.worker_rank(0), group=wgroup)
            eng.refresh_compute_weights()
            if hasattr(comm, "seed_counters"):     # async mode keeps global_step / Adam t in the shard counters
                comm.seed_counters(eng)

    if FLAGS.log_device_placement and is_chief:
        log_placement(trainer, role, use_cuda)
    replica = 0 if FLAGS.identical_replicas else role.task_index
    # --gpu_decode: the training pipeline ships the float64 records as stored and casts them on the device
    images = _make_reader(FLAGS, FLAGS.data_dir, batch, cfg, FLAGS.seed + 17 * replica,
                          raw_f64=bool(FLAGS.gpu_decode) and use_cuda and not FLAGS.synthetic_data)
    sample_images = None
    if is_chief:
        # the second pipeline (image_train.py:84) is built before the loop, as in the reference graph: filling its
        # shuffle queue inside the step loop would keep the other replicas waiting at their next exchange
        sample_dir = FLAGS.sample_image_dir if os.path.isdir(FLAGS.sample_image_dir) else FLAGS.data_dir
        sample_images = _make_reader(FLAGS, sample_dir, batch, cfg, FLAGS.seed + 999)
    rng = np.random.RandomState(FLAGS.seed + 1000 + replica)
    sample_z = rng.uniform(-1, 1, size=(SAMPLE_SIZE, cfg.z_dim)).astype(np.float32)  # image_train.py:77
    prefetch = None
    if use_cuda:
        from .data.loader import DevicePrefetcher
        prefetch = DevicePrefetcher(images, batch, (cfg.output_size, cfg.output_size, cfg.c_dim), trainer.device,
                                    gpu_decode=getattr(images, "raw_f64", False))

    step = trainer.global_step
    start_time = time.time()
    step_times = []
    local_iters = 0
    while sv.keep_going(step):
        batch_z = rng.uniform(-1, 1, [batch, cfg.z_dim]).astype(np.float32)        # image_train.py:151-152
        if prefetch is not None:
            real_img = prefetch.next()
        else:
            real_img = images.dequeue_many(batch)                                   # sess.run(images), :153
            if real_img.shape[0] < batch:
                break
        want_summary = sv.summary_due()
        t0 = time.time()
        d_real, d_fake, g_loss, step = trainer.train_step(real_img, batch_z, record=want_summary)
        step_times.append(time.time() - t0)
        local_iters += 1
        if FLAGS.log_every and local_iters % FLAGS.log_every == 0:
            print("Epoch: [%2d] step: [%2d] time: %4.4f, d_loss: %.8f, g_loss: %.8f"
                  % (step // spe, step % spe, time.time() - start_time, d_fake + d_real, g_loss))
        if want_summary:
            summ = trainer.summaries(real_img, batch_z, (d_real, d_fake, g_loss)) if is_chief else b""
            sv.summary_computed(summ, step)
        if prefetch is not None:
            prefetch.release()
        if step % FLAGS.sample_every == 1 and is_chief:                               # image_train.py:179
            sample_image = sample_images.dequeue_many(batch)
            grid = trainer.sample_grid(sample_z[:batch])
            d1_loss, g1_loss = trainer.evaluate_losses(sample_image, sample_z[:batch])
            S.save_merged(grid, S.sample_path(FLAGS.sample_dir, step, spe))
            print("[Sample] d_loss: %.8f, g_loss: %.8f" % (d1_loss, g1_loss))
        sv.maybe_checkpoint(trainer, step)

    # ---- orderly shutdown -------------------------------------------------------------------------------------------
    sv.maybe_checkpoint(trainer, step, force=True)
    if FLAGS.profile and step_times:
        st = np.array(step_times[min(3, len(step_times) - 1):])
        print(f"[profile] {len(st)} steps: mean {st.mean()*1e3:.3f} ms, p50 {np.median(st)*1e3:.3f} ms, "
              f"{batch / st.mean():.1f} images/s per worker")
    if ps_client is not None:
        ps_client.done()
    if prefetch is not None:
        prefetch.close()
    sv.close()
    if role.distributed:
        if use_cuda and comm is not None and hasattr(comm, "close"):
            comm.close()
        if use_cuda and role.cluster.num_ps:
            import torch.distributed as dist
            dist.barrier()   # pairs with run_ps: workers finished
        cl.shutdown()
    return 0


def log_placement(trainer, role, use_cuda: bool) -> None:
    """``--log_device_placement`` (image_train.py:36; the reference declares it and never reads it): one line per
    variable with the device -- and, for sharded multi-GPU state, the ranks -- that hold it."""
    eng = getattr(trainer, "engine", None)
    if eng is None:
        where = "ps shards (parallel/ps.py)" if role.cluster.num_ps else "cpu"
        for name in trainer.params:
            print(f"{name}: {where}")
        return
    comm, A = eng.comm, eng.arena
    for name, (off, shape) in A.offsets.items():
        n = 1
        for d in shape:
            n *= d
        if comm is not None and hasattr(comm, "owner_ranks"):
            owners = comm.owner_ranks(off, off + n)
            print(f"{name}: {eng.device} (bf16 copy on every rank; f32 master + Adam slots on rank(s) {owners})")
        else:
            print(f"{name}: {eng.device}")


def _worker_barrier(role):
    """Barrier among workers only (ps tasks are busy serving): workers wait for the chief's init."""
    import torch.distributed as dist
    ranks = [role.cluster.worker_rank(i) for i in range(role.cluster.num_workers)]
    # new_group is collective over the whole world in torch; ps tasks never call it, so use p2p tokens
    chief = ranks[0]
    token = torch.zeros(1)
    if role.rank == chief:
        for r in ranks[1:]:
            dist.send(token, r)
    else:
        dist.recv(token, chief)
    return None


def train(FLAGS) -> int:
    cfg = F.model_config_from_flags()
    role = cl.resolve_role(FLAGS.job_name, FLAGS.task_index, FLAGS.ps_hosts, FLAGS.worker_hosts)
    if role.is_ps:
        return run_ps(FLAGS, role, cfg)
    return run_worker(FLAGS, role, cfg)


def main(argv=None) -> int:
    from absl import flags as absl_flags
    F.define_flags()
    FLAGS = absl_flags.FLAGS
    FLAGS(argv if argv is not None else sys.argv)
    F.pprint_flags()                                      # image_train.py:223
    os.makedirs(FLAGS.checkpoint_dir, exist_ok=True)      # :225-226
    os.makedirs(FLAGS.sample_dir, exist_ok=True)          # :227-228
    return train(FLAGS)


if __name__ == "__main__":
    sys.exit(main())
