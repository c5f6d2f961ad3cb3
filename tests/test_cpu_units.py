"""CPU-only unit tests: oracle geometry, manifest / checkpoint contract, input pipeline, summaries,
sample grid, flags and cluster bootstrap."""
import math
import os

import numpy as np
import pytest
import torch

from distributed_tensorflow_for_dcgan_b200.config import ModelConfig, steps_per_epoch, WIDE_128
from distributed_tensorflow_for_dcgan_b200.models import dcgan, manifest
from distributed_tensorflow_for_dcgan_b200.ops import functional as Fn


# ---- geometry ---------------------------------------------------------------------
def test_same_padding_is_one_before_two_after():
    assert Fn.same_padding(64) == (32, 1, 2)
    assert Fn.same_padding(8) == (4, 1, 2)
    assert Fn.same_padding(7) == (4, 2, 2)


def test_conv_same_is_not_symmetric_padding_and_deconv_is_its_gradient():
    torch.manual_seed(0)
    x = torch.randn(2, 8, 8, 3, dtype=torch.float64, requires_grad=True)
    w = torch.randn(5, 5, 3, 4, dtype=torch.float64)
    y = Fn.conv2d_same(x, w)
    assert y.shape == (2, 4, 4, 4)
    sym = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), stride=2, padding=2)
    assert not torch.allclose(y.permute(0, 3, 1, 2), sym)
    gy = torch.randn_like(y)
    gx, = torch.autograd.grad(y, x, gy)
    assert torch.allclose(gx, Fn.deconv2d(gy, w))
    # and PyTorch's padding=2/output_padding=1 transposed conv is the mirror image, not this
    pt = torch.nn.functional.conv_transpose2d(gy.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), stride=2, padding=2,
                                              output_padding=1).permute(0, 2, 3, 1)
    assert not torch.allclose(gx, pt)


def test_parity_tap_counts():
    taps = {(ph, pw): sum(1 for kh in range(5) for kw in range(5) if (ph + 1 - kh) % 2 == 0 and (pw + 1 - kw) % 2 == 0)
            for ph in (0, 1) for pw in (0, 1)}
    assert sorted(taps.values()) == [4, 6, 6, 9]


# ---- manifest ----------------------------------------------------------------------------
def test_parameter_counts_and_names():
    cfg = ModelConfig()
    c = manifest.param_counts(cfg)
    assert c == {"G": 5135363, "D": 4316545, "total": 9451908}
    names = [s.name for s in manifest.trainable_specs(cfg)]
    assert len(names) == 34 and len(set(names)) == 34
    shapes = {s.name: s.shape for s in manifest.trainable_specs(cfg)}
    assert shapes["g_h0_lin/Matrix"] == (100, 8192) and shapes["g_h1/w"] == (5, 5, 256, 512)
    assert shapes["g_h4/w"] == (5, 5, 3, 64) and shapes["d_h0_conv/w"] == (5, 5, 3, 64)
    assert shapes["d_h3_conv/w"] == (5, 5, 256, 512) and shapes["d_h3_lin/Matrix"] == (8192, 1)
    assert "d_bn0/beta" not in shapes  # d_bn0 is never applied (distriubted_model.py:55,118)
    d = [n for n in names if manifest.is_d_var(n)]
    g = [n for n in names if manifest.is_g_var(n)]
    assert set(d).isdisjoint(g) and len(d) == 16 and len(g) == 18
    assert len(manifest.ema_specs(cfg)) == 20
    assert steps_per_epoch(64) == 5051


def test_wide_config_shapes():
    p = dcgan.init_params(WIDE_128)
    z = torch.rand(2, 100) * 2 - 1
    G = dcgan.generator(z, p, WIDE_128)
    assert G.shape == (2, 128, 128, 3)
    _, l = dcgan.discriminator(G, p, WIDE_128)
    assert l.shape == (2, 1)


def test_initialisers():
    p = dcgan.init_params(ModelConfig(), seed=1)
    assert abs(p["g_h1/w"].std().item() - 0.02) < 1e-3
    assert p["d_h2_conv/w"].abs().max().item() <= 0.04 + 1e-6  # truncated at two sigma
    assert abs(p["g_bn1/gamma"].mean().item() - 1.0) < 5e-3 and p["g_bn1/beta"].abs().max() == 0
    assert p["g_h0_lin/bias"].abs().max() == 0


# ---- ops -----------------------------------------------------------------------------------------------------
def test_batch_norm_and_ema():
    x = torch.randn(4, 8, 8, 16) * 3 + 1
    m, v = Fn.batch_moments(x)
    assert torch.allclose(v, x.reshape(-1, 16).var(0, unbiased=False), atol=1e-5)
    y = Fn.batch_norm_apply(x, m, v, torch.zeros(16), torch.ones(16))
    assert y.reshape(-1, 16).mean(0).abs().max() < 1e-5
    shadow = torch.zeros(16)
    Fn.ema_update_(shadow, m, 0.9)
    assert torch.allclose(shadow, 0.1 * m)


def test_bce_matches_torch():
    x = torch.randn(64, 1) * 3
    for t in (0.0, 1.0):
        ref = torch.nn.functional.binary_cross_entropy_with_logits(x, torch.full_like(x, t))
        assert torch.allclose(Fn.bce_with_logits_mean(x, t), ref, atol=1e-6)


def test_adam_tf_form():
    from distributed_tensorflow_for_dcgan_b200.optim.adam_tf import adam_tf_step_
    p, g = torch.tensor([1.0], dtype=torch.float64), torch.tensor([0.5], dtype=torch.float64)
    m, v = torch.zeros(1, dtype=torch.float64), torch.zeros(1, dtype=torch.float64)
    adam_tf_step_(p, g, m, v, 1)
    lr_t = 2e-4 * math.sqrt(1 - 0.999) / (1 - 0.5)
    exp = 1.0 - lr_t * (0.5 * 0.5) / (math.sqrt(0.001 * 0.25) + 1e-8)
    assert abs(p.item() - exp) < 1e-12
    # differs from torch.optim.Adam, which adds eps to the bias-corrected sqrt(v)
    q = torch.nn.Parameter(torch.tensor([1.0], dtype=torch.float64))
    q.grad = torch.tensor([1e-7], dtype=torch.float64)   # small gradient: eps placement matters
    torch.optim.Adam([q], lr=2e-4, betas=(0.5, 0.999), eps=1e-8).step()
    p2 = torch.tensor([1.0], dtype=torch.float64)
    adam_tf_step_(p2, torch.tensor([1e-7], dtype=torch.float64), torch.zeros(1, dtype=torch.float64),
                  torch.zeros(1, dtype=torch.float64), 1)
    assert abs(q.item() - p2.item()) > 1e-6


def test_simultaneous_update_semantics():
    """D and G gradients come from ONE forward pass; global_step advances once (image_train.py:156-158)."""
    from distributed_tensorflow_for_dcgan_b200.trainer import CPUTrainer
    cfg = ModelConfig(batch_size=2)
    tr = CPUTrainer(cfg, seed=0)
    before = {k: v.clone() for k, v in tr.params.items()}
    real = torch.rand(2, 64, 64, 3) * 2 - 1
    z = torch.rand(2, 100) * 2 - 1
    out = tr.train_step(real, z)
    assert out[3] == 1
    moved = [k for k in before if not torch.equal(before[k], tr.params[k])]
    assert len(moved) >= 27
    ema = tr.ema["g_bn0/moments/mean/ExponentialMovingAverage"]
    assert ema.abs().max() > 0 and tr.ema["d_bn1_1/moments/variance/ExponentialMovingAverage"].abs().max() > 0
    imgs = tr.sample(z)
    assert imgs.shape == (2, 64, 64, 3)


# ---- checkpoint ------------------------------------------------------------------------------------------------
def test_checkpoint_roundtrip_and_resume(tmp_path):
    from distributed_tensorflow_for_dcgan_b200.trainer import CPUTrainer
    from distributed_tensorflow_for_dcgan_b200.utils import checkpoint
    cfg = ModelConfig(batch_size=2)
    tr = CPUTrainer(cfg, seed=0)
    real, z = torch.rand(2, 64, 64, 3) * 2 - 1, torch.rand(2, 100) * 2 - 1
    tr.train_step(real, z)
    path = checkpoint.save(str(tmp_path), tr.state_dict(), tr.global_step)
    assert os.path.basename(path) == "model.ckpt-1"
    text = open(tmp_path / "checkpoint").read()
    assert 'model_checkpoint_path: "model.ckpt-1"' in text
    st = checkpoint.get_checkpoint_state(str(tmp_path))
    assert st.model_checkpoint_path.endswith("model.ckpt-1")
    tr2 = CPUTrainer(cfg, seed=5)
    tr2.load_state_dict(checkpoint.load_latest(str(tmp_path)))
    assert tr2.global_step == 1 and tr2.opt_d.t == 1
    a = tr.train_step(real, z)
    b = tr2.train_step(real, z)
    assert a[3] == b[3] == 2 and abs(a[0] - b[0]) < 1e-6 and abs(a[2] - b[2]) < 1e-6
    for k in tr.params:
        assert torch.allclose(tr.params[k], tr2.params[k], atol=1e-7), k
    assert checkpoint.get_checkpoint_state(str(tmp_path / "nope")) is None


# ---- input pipeline -----------------------------------------------------------------------------------------------
def _write_dataset(tmp_path, n_files=3, per_file=40):
    from distributed_tensorflow_for_dcgan_b200.data import tfrecord
    rng = np.random.RandomState(0)
    imgs = rng.rand(n_files * per_file, 64, 64, 3) * 2 - 1
    d = tmp_path / "train"
    d.mkdir()
    for i in range(n_files):
        tfrecord.write_image_tfrecord(str(d / f"image_{i}.tfrecords"), imgs[i * per_file:(i + 1) * per_file])
    return str(d), imgs


def test_tfrecord_roundtrip_bit_exact_and_crc(tmp_path):
    from distributed_tensorflow_for_dcgan_b200.data import tfrecord
    from distributed_tensorflow_for_dcgan_b200.utils.crc32c import crc32c_py
    assert crc32c_py(b"123456789") == 0xE3069283
    d, imgs = _write_dataset(tmp_path, 1, 5)
    f = os.path.join(d, "image_0.tfrecords")
    assert os.path.getsize(f) == 491755  # 5 x (98,304 B payload + 16 B framing + 31 B Example envelope)
    out = [tfrecord.decode_image_example(r) for r in tfrecord.read_records(f)]
    assert len(out) == 5 and all(o.dtype == np.float32 for o in out)
    assert np.array_equal(out[2], imgs[2].astype(np.float32))
    raw = bytearray(open(f, "rb").read())
    raw[100] ^= 0xFF
    open(f, "wb").write(raw)
    with pytest.raises(IOError):
        list(tfrecord.read_records(f))


def test_native_reader_matches_python_reader(tmp_path):
    from distributed_tensorflow_for_dcgan_b200.data import loader, native_reader, tfrecord
    d, imgs = _write_dataset(tmp_path)
    files = tfrecord.list_data_files(d)
    nat = native_reader.NativeShuffleReader(files, capacity=60, min_after_dequeue=30, num_threads=4, seed=3, epochs=1)
    got = []
    while True:
        b = nat.dequeue_many(16)
        if b.shape[0] == 0:
            break
        got.append(b.copy())
    nat.close()
    got = np.concatenate(got)
    ref = imgs.astype(np.float32)
    assert sorted(a.tobytes() for a in got) == sorted(a.tobytes() for a in ref)  # every record once, bit exact
    assert not np.array_equal(got, ref)                                          # and shuffled
    py = loader.PythonShuffleReader(files, capacity=60, min_after_dequeue=30, seed=3, epochs=1)
    n = 0
    while True:
        b = py.dequeue_many(16)
        if b.shape[0] == 0:
            break
        n += b.shape[0]
    assert n == ref.shape[0]


def test_native_reader_raw_float64_mode(tmp_path):
    """raw_f64: the queue hands out the records as stored (float64, no cast on the reader threads); the float32 cast
    of the default mode equals casting these afterwards -- which is what the GPU decode path does on the device."""
    from distributed_tensorflow_for_dcgan_b200.data import native_reader, tfrecord
    d, imgs = _write_dataset(tmp_path, 2, 20)
    files = tfrecord.list_data_files(d)
    raw = native_reader.NativeShuffleReader(files, capacity=30, min_after_dequeue=10, num_threads=2, seed=5, epochs=1,
                                            raw_f64=True)
    got = []
    while True:
        b = raw.dequeue_many(8)
        if b.shape[0] == 0:
            break
        assert b.dtype == np.float64
        got.append(b.copy())
    raw.close()
    got = np.concatenate(got)
    assert sorted(a.tobytes() for a in got) == sorted(a.tobytes() for a in imgs)   # bit-exact float64 payloads
    assert sorted(a.astype(np.float32).tobytes() for a in got) == sorted(a.astype(np.float32).tobytes() for a in imgs)


def test_shuffle_queue_waits_for_min_after_dequeue(tmp_path):
    from distributed_tensorflow_for_dcgan_b200.data import loader, native_reader, tfrecord
    assert loader.queue_sizes(64) == (10776, 10968)  # image_input.py:134-136,82-84
    d, _ = _write_dataset(tmp_path, 2, 40)
    r = native_reader.NativeShuffleReader(tfrecord.list_data_files(d), capacity=50, min_after_dequeue=40,
                                          num_threads=2, seed=0)  # endless
    b = r.dequeue_many(8)
    assert b.shape == (8, 64, 64, 3)
    assert r.size() >= 40 - 8   # a dequeue only happens while more than min_after_dequeue elements are queued
    r.close()


def test_distorted_inputs_api(tmp_path, capsys):
    from distributed_tensorflow_for_dcgan_b200.data import loader
    d, _ = _write_dataset(tmp_path, 2, 20)
    r = loader.distorted_inputs(d, batch_size=4, min_after_dequeue=8, capacity=20, num_threads=2)
    assert "Filling queue with 8 CIFAR images" in capsys.readouterr().out
    assert r.dequeue_many(4).shape == (4, 64, 64, 3)
    r.close()
    with pytest.raises((ValueError, FileNotFoundError)):
        loader.distorted_inputs(str(tmp_path / "missing"))


# ---- summaries / samples -------------------------------------------------------------------------------------------
def test_summary_events_and_sample_grid(tmp_path):
    from distributed_tensorflow_for_dcgan_b200.trainer import CPUTrainer
    from distributed_tensorflow_for_dcgan_b200.utils import sample, summary
    cfg = ModelConfig(batch_size=2)
    tr = CPUTrainer(cfg)
    real, z = torch.rand(2, 64, 64, 3) * 2 - 1, torch.rand(2, 100) * 2 - 1
    out = tr.train_step(real, z, record=True)
    w = summary.SummaryWriter(str(tmp_path))
    w.add_summary(tr.summaries(real, z, out[:3]), out[3])
    w.close()
    events = list(summary.read_events(w.path))
    step, tags = events[-1]
    assert step == 1
    for t in ("z", "d", "d_", "G/image/0", "d_loss_real", "d_loss_fake", "g_loss", "d_loss", "g_h1/w",
              "g_h0_relu/activations", "d_h2_conv/sparsity", "d_h3_lin_1/activations"):
        assert t in tags, t
    assert tags["d_loss"][0] == "scalar" and abs(tags["d_loss"][1] - (out[0] + out[1])) < 1e-5
    imgs = np.stack([np.full((4, 4, 3), i, dtype=np.float64) for i in range(6)])
    grid = sample.merge(imgs, [2, 3])
    assert grid.shape == (8, 12, 3) and grid[0, 4, 0] == 1 and grid[4, 0, 0] == 3  # row-major tiling
    assert np.allclose(sample.inverse_transform(np.array([-1.0, 1.0])), [0.0, 1.0])
    assert sample.sample_path("s", 5052, 5051).endswith("train_01_0001.png")
    sample.save_images(np.random.rand(64, 8, 8, 3) * 2 - 1, [8, 8], str(tmp_path / "g.png"))
    from PIL import Image
    assert Image.open(tmp_path / "g.png").size == (64, 64)


# ---- flags / cluster -----------------------------------------------------------------------------------------------
def test_reference_flag_names_and_defaults():
    from absl import flags as absl_flags
    from distributed_tensorflow_for_dcgan_b200 import flags as F
    F.define_flags()
    fv = absl_flags.FLAGS
    if not fv.is_parsed():
        fv(["test"])
    expect = {"epoch": 25, "learning_rate": 0.0002, "beta1": 0.5, "batch_size": 64, "image_size": 108,
              "output_size": 64, "c_dim": 3, "dataset": "celebA", "checkpoint_dir": "checkpoint",
              "sample_dir": "samples", "is_train": False, "is_crop": False, "visualize": False, "data_dir": "train",
              "sample_image_dir": "sample_data", "ps_hosts": "", "worker_hosts": "", "job_name": "", "task_index": 0,
              "log_device_placement": True, "save_summaries_secs": 10}
    for k, v in expect.items():
        assert fv[k].default == v, k
    assert math.isinf(fv["train_size"].default)
    for k in ("sync_replicas", "update_mode", "dtype", "max_steps", "synthetic_data", "ngf", "ndf", "z_dim", "sync_bn",
              "comm_backend"):
        assert k in fv


def test_cluster_spec_and_roles():
    from distributed_tensorflow_for_dcgan_b200.parallel import cluster as cl
    spec = cl.ClusterSpec.from_flags("a:1,b:2", "c:3,d:4,e:5")
    assert spec.num_ps == 2 and spec.num_workers == 3 and spec.world_size == 5
    assert spec.global_rank("ps", 1) == 1 and spec.global_rank("worker", 0) == 2
    assert spec.master_endpoint() == ("a", 1)
    r = cl.resolve_role("worker", 0, "a:1", "c:3,d:4")
    assert r.is_chief and not r.is_ps and r.rank == 1 and r.world_size == 3
    assert cl.resolve_role("ps", 0, "a:1", "c:3").is_ps
    assert cl.resolve_role("worker", 1, "a:1", "c:3,d:4").is_chief is False
    solo = cl.resolve_role("", 0, "", "")
    assert solo.is_chief and not solo.distributed
    with pytest.raises(ValueError):
        spec.global_rank("worker", 7)


def test_variable_sharding_is_balanced_and_exhaustive():
    from distributed_tensorflow_for_dcgan_b200.parallel.ps import shard_variables
    specs = manifest.trainable_specs(ModelConfig())
    shards = shard_variables(specs, 4)
    flat = [n for s in shards for n in s]
    assert sorted(flat) == sorted(s.name for s in specs)
    sizes = [sum(next(x.numel for x in specs if x.name == n) for n in s) for s in shards]
    assert max(sizes) <= 3276800 + 100000  # no shard much beyond the largest single tensor


def test_worker_group_ranks_exclude_ps_tasks():
    """The workers' own process group (symmetric-memory rendezvous, initial broadcast) is made of the worker ranks
    only; ps tasks come first in the rank order and stay out of it (reference launch: image_train.py:52-67)."""
    from distributed_tensorflow_for_dcgan_b200.parallel import cluster as cl
    spec = cl.ClusterSpec.from_flags("a:1,b:2", "c:3,d:4,e:5")
    assert cl.worker_ranks(spec) == [2, 3, 4]
    assert cl.worker_ranks(cl.ClusterSpec.from_flags("", "c:3,d:4")) == [0, 1]
    solo = cl.resolve_role("", 0, "", "")
    assert cl.worker_group(solo, "gloo") is None   # a single process has no group to build


def test_train_size_flag_shortens_the_epoch():
    from distributed_tensorflow_for_dcgan_b200.config import steps_per_epoch
    assert steps_per_epoch(64) == 5051                      # image_train.py:44-48 under Python 2
    assert steps_per_epoch(64, float("inf")) == 5051
    assert steps_per_epoch(64, 6400) == 100
    assert steps_per_epoch(64, 10) == 1                     # never zero: the log line divides by it
    assert steps_per_epoch(64, 10 ** 9) == 5051


def test_flag_docstring_matches_what_is_read():
    """The flags module says which reference-dead flags are live; hold it to that."""
    import inspect
    from distributed_tensorflow_for_dcgan_b200 import flags as F, train as T
    src = inspect.getsource(T)
    for live in ("train_size", "log_device_placement", "sample_dir"):
        assert f"FLAGS.{live}" in src, live
    for unused in F.UNUSED_IN_REFERENCE_AND_HERE:
        assert f"FLAGS.{unused}" not in src, unused
