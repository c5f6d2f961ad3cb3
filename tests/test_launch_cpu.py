"""Launch surface on CPU + gloo ("Launch surface", "Distributed correctness (CPU)").
BASELINE.json config 1: DCGAN 64x64 synthetic images, batch 16, 1 ps + 2 workers, sync mode."""
import os
import socket
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_ports(n):
    socks, ports = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    return ports


def _launch(job, idx, ps_hosts, worker_hosts, extra, tmp):
    cmd = [sys.executable, os.path.join(ROOT, "image_train.py"), f"--job_name={job}", f"--task_index={idx}",
           f"--ps_hosts={ps_hosts}", f"--worker_hosts={worker_hosts}", "--device=cpu", "--synthetic_data",
           f"--checkpoint_dir={tmp}/ckpt", f"--sample_dir={tmp}/samples", "--sample_every=1000000"] + extra
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2")
    return subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, cwd=str(tmp))


def _run_cluster(tmp, n_ps, n_workers, extra):
    ports = _free_ports(n_ps + n_workers)
    ps_hosts = ",".join(f"127.0.0.1:{p}" for p in ports[:n_ps])
    worker_hosts = ",".join(f"127.0.0.1:{p}" for p in ports[n_ps:])
    procs = [("ps", i, _launch("ps", i, ps_hosts, worker_hosts, extra, tmp)) for i in range(n_ps)]
    procs += [("worker", i, _launch("worker", i, ps_hosts, worker_hosts, extra, tmp)) for i in range(n_workers)]
    outs = {}
    for job, i, p in procs:
        try:
            out, _ = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for _, _, q in procs:
                q.kill()
            raise
        outs[(job, i)] = out
        assert p.returncode == 0, f"{job} {i} failed:\n{out[-3000:]}"
    return outs


def test_one_ps_two_workers_sync(tmp_path):
    outs = _run_cluster(tmp_path, 1, 2, ["--batch_size=16", "--max_steps=3", "--sync_replicas=true"])
    # ps joined, served three synchronous rounds and left cleanly once both workers were done
    assert "served 3 applies, global_step 3" in outs[("ps", 0)]
    lines = {k: [l for l in v.splitlines() if l.startswith("Epoch:")] for k, v in outs.items() if k[0] == "worker"}
    assert len(lines[("worker", 0)]) == 3 and len(lines[("worker", 1)]) == 3
    assert "step: [ 3]" in lines[("worker", 0)][-1] and "step: [ 3]" in lines[("worker", 1)][-1]
    # the chief (worker 0) wrote the checkpoint; it carries the TF names and the ps-side Adam slots
    from distributed_tensorflow_for_dcgan_b200.utils import checkpoint
    sd = checkpoint.load_latest(str(tmp_path / "ckpt"))
    assert sd is not None and int(sd["Variable"]) == 3
    assert sd["g_h1/w"].shape == (5, 5, 256, 512) and sd["d_h1_conv/w"].shape == (5, 5, 64, 128)
    assert sd["g_h1/w/Adam"].abs().max() > 0


def test_two_ps_two_workers_async(tmp_path):
    outs = _run_cluster(tmp_path, 2, 2, ["--batch_size=4", "--max_steps=6", "--sync_replicas=false"])
    # async: global_step counts the applies of *all* workers; nothing is lost
    served = [int(outs[("ps", i)].split("served ")[1].split(" applies")[0]) for i in range(2)]
    steps = sum(len([l for l in outs[("worker", i)].splitlines() if l.startswith("Epoch:")]) for i in range(2))
    assert served[0] == served[1] == steps and steps >= 6


def test_ps_less_sync_workers_stay_identical(tmp_path):
    outs = _run_cluster(tmp_path, 0, 2, ["--batch_size=4", "--max_steps=2"])
    l0 = [l for l in outs[("worker", 0)].splitlines() if l.startswith("Epoch:")]
    assert len(l0) == 2


def test_sync_replicas_reproduce_single_process_run(tmp_path):
    """Distributed correctness (CPU): with every replica fed the same batch, the mean gradient equals
    the single-process gradient, so 1 ps + 2 workers (sync) must land on the same parameters."""
    from distributed_tensorflow_for_dcgan_b200.utils import checkpoint
    a, b = tmp_path / "dist", tmp_path / "solo"
    a.mkdir(); b.mkdir()
    _run_cluster(a, 1, 2, ["--batch_size=4", "--max_steps=2", "--identical_replicas"])
    p = _launch("worker", 0, "", "", ["--batch_size=4", "--max_steps=2"], b)
    out, _ = p.communicate(timeout=600)
    assert p.returncode == 0, out[-2000:]
    sd_a = checkpoint.load_latest(str(a / "ckpt"))
    sd_b = checkpoint.load_latest(str(b / "ckpt"))
    for k in ("g_h1/w", "d_h2_conv/w", "g_bn1/gamma", "d_h3_lin/Matrix", "g_h2/w/Adam_1"):
        assert torch.allclose(sd_a[k], sd_b[k], rtol=1e-4, atol=1e-6), k


def _kill_worker_one(tmp_path, mode_flags, budget):
    import threading
    import time
    ports = _free_ports(3)
    ps_hosts = f"127.0.0.1:{ports[0]}"
    worker_hosts = ",".join(f"127.0.0.1:{p}" for p in ports[1:])
    extra = ["--batch_size=4", f"--max_steps={budget}", "--worker_timeout_secs=6"] + mode_flags
    procs = {"ps": _launch("ps", 0, ps_hosts, worker_hosts, extra, tmp_path),
             "w0": _launch("worker", 0, ps_hosts, worker_hosts, extra, tmp_path),
             "w1": _launch("worker", 1, ps_hosts, worker_hosts, extra, tmp_path)}
    lines = {k: [] for k in procs}

    def pump(tag):
        for l in procs[tag].stdout:
            lines[tag].append(l.rstrip())
    threads = [threading.Thread(target=pump, args=(k,), daemon=True) for k in procs]
    for t in threads:
        t.start()
    try:
        t0 = time.time()
        while time.time() - t0 < 300 and sum(l.startswith("Epoch:") for l in lines["w1"]) < 2:
            assert procs["w1"].poll() is None, "\n".join(lines["w1"][-20:])
            time.sleep(0.2)
        procs["w1"].kill()                               # the non-chief worker dies mid-run
        assert procs["w0"].wait(timeout=300) == 0, "\n".join(lines["w0"][-20:])
        assert procs["ps"].wait(timeout=120) == 0, "\n".join(lines["ps"][-20:])
    finally:
        for p in procs.values():
            if p.poll() is None:
                p.kill()
    for t in threads:
        t.join(timeout=5)
    return lines


def test_async_mode_survives_a_killed_worker(tmp_path):
    """Fault tier of SURVEY.md 4 / 5.3: asynchronous replicas tolerate the loss of a non-chief worker by
    construction (the others keep training); here the parameter server additionally notices the silent worker
    at shutdown and leaves instead of blocking forever like ``server.join()`` (image_train.py:62-63)."""
    lines = _kill_worker_one(tmp_path, ["--sync_replicas=false"], 16)
    assert any("step: [16]" in l for l in lines["w0"])   # the survivor reached the step budget
    assert any("presumed dead and dropped" in l for l in lines["ps"])
    assert any("served" in l and "applies" in l for l in lines["ps"])


def test_sync_mode_drops_a_dead_worker_and_continues(tmp_path):
    """Synchronous rounds wait for every live worker; one that has been silent for --worker_timeout_secs while
    a round is open is dropped and the round completes with the survivors (the reference would hang)."""
    lines = _kill_worker_one(tmp_path, ["--sync_replicas=true"], 6)
    assert any("step: [ 6]" in l for l in lines["w0"])
    assert any("presumed dead and dropped" in l for l in lines["ps"])
