"""Distributed correctness on GPUs: fused in-kernel reduce-scatter + Adam + all-gather ==
NCCL all-reduce + Adam (P2P and NVLS multicast paths), cross-GPU batch-norm statistics, async
parameter-server mode.  Needs >= 2 GPUs; launches scripts/multigpu_check.py under torchrun."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_fused_comm_two_ranks(cuda_lib):
    n = min(torch.cuda.device_count(), 8)
    n = 2 if n < 4 else (4 if n < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "scripts", "multigpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0 and "MULTIGPU CHECK PASSED" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def _free_ports(n):
    import socket
    socks, ports = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    return ports


def _launch_cluster(tmp_path, n_ps, n_workers, extra, timeout=600):
    """The reference's launch surface (image_train.py:28-35) on GPUs: one process per task."""
    ports = _free_ports(n_ps + n_workers)
    ps_hosts = ",".join(f"127.0.0.1:{p}" for p in ports[:n_ps])
    worker_hosts = ",".join(f"127.0.0.1:{p}" for p in ports[n_ps:])
    procs = []
    for job, count in (("ps", n_ps), ("worker", n_workers)):
        for i in range(count):
            env = dict(os.environ, PYTHONPATH=ROOT)
            if job == "worker":
                env["LOCAL_RANK"] = str(i)
            else:
                env["CUDA_VISIBLE_DEVICES"] = ""   # a ps task is a CPU-only joiner on the GPU path
            cmd = [sys.executable, os.path.join(ROOT, "image_train.py"), f"--job_name={job}", f"--task_index={i}",
                   f"--ps_hosts={ps_hosts}", f"--worker_hosts={worker_hosts}", "--device=cuda", "--synthetic_data",
                   f"--checkpoint_dir={tmp_path}/ckpt", f"--sample_dir={tmp_path}/samples",
                   "--save_summaries_secs=0", "--nolog_device_placement"] + extra
            procs.append((job, i, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                                                   cwd=str(tmp_path), env=env)))
    outs = []
    for job, i, p in procs:
        try:
            out, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            p.kill()
            out, _ = p.communicate()
            out += "\n[TIMEOUT]"
        outs.append((job, i, p.returncode, out))
    return outs


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_ps_plus_two_gpu_workers(cuda_lib, tmp_path):
    """--job_name=ps next to GPU workers: the ps task joins (the shards live on the workers' GPUs) and leaves
    when the workers are done; all three processes exit 0."""
    outs = _launch_cluster(tmp_path, 1, 2, ["--max_steps=5"])
    for job, i, rc, out in outs:
        assert rc == 0, f"{job} {i} rc={rc}\n{out[-3000:]}"
    chief = [o for o in outs if o[0] == "worker" and o[1] == 0][0][3]
    assert "step: [ 5]" in chief or "step: [5]" in chief, chief[-2000:]
    assert any("joined" in o[3] for o in outs if o[0] == "ps")


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_sync_bn_with_chief_only_sampling(cuda_lib, tmp_path):
    """Cross-replica batch-norm statistics are a collective; the chief's sample-time evaluation
    (image_train.py:179-184) runs on one rank only and must stay out of it.  sample_every=2 makes the chief
    evaluate three times within six steps; every step after that still has to complete on both ranks."""
    outs = _launch_cluster(tmp_path, 0, 2, ["--max_steps=6", "--sync_bn", "--sample_every=2"])
    for job, i, rc, out in outs:
        assert rc == 0, f"{job} {i} rc={rc}\n{out[-3000:]}"
        assert "error flag" not in out
    chief = outs[0][3]
    assert chief.count("[Sample] d_loss:") >= 2, chief[-2000:]
