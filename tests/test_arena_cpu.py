"""Layout of the flat parameter arena (engine.ParamArena) -- the contract the optimizer / communication kernels
rely on: [G bf16-consumed | G f32-consumed | D bf16-consumed | D f32-consumed], every tensor 32-byte aligned,
TF names and TF layouts, padded to whole 1024-element blocks."""
import pytest
import torch

from distributed_tensorflow_for_dcgan_b200.config import ModelConfig
from distributed_tensorflow_for_dcgan_b200.engine import ParamArena
from distributed_tensorflow_for_dcgan_b200.models import manifest


def test_arena_layout():
    cfg = ModelConfig(batch_size=4)
    A = ParamArena(cfg, "cpu")
    specs = {s.name: s for s in manifest.trainable_specs(cfg)}
    assert set(A.offsets) == set(specs) and len(specs) == 34
    assert sum(s.numel for s in specs.values()) == 9451908      # G 5,135,363 + D 4,316,545
    spans = sorted((off, off + specs[n].numel, n) for n, (off, _) in A.offsets.items())
    for (a0, a1, an), (b0, _, bn) in zip(spans, spans[1:]):
        assert a1 <= b0, (an, bn)                                # no overlap
    for off, _, name in spans:
        assert off % 8 == 0, name                                # 32-byte aligned f32 (16-byte aligned bf16)
    assert A.padded % 1024 == 0 and A.numel <= A.padded < A.numel + 1024
    # region order
    for name, (off, _) in A.offsets.items():
        is_g = manifest.is_g_var(name)
        assert (off < A.g_end) == is_g, name
        key = "g" if is_g else "d"
        assert (off < A.f32_start[key]) == A.reads_bf16(name), name
    assert A.f32_start["g"] < A.g_end < A.f32_start["d"] < A.numel
    # the tensors the kernels read in f32: biases, batch-norm parameters, the 3-channel filters, d_h3_lin
    f32_names = {n for n in specs if not A.reads_bf16(n)}
    assert {"g_h4/w", "d_h0_conv/w", "d_h3_lin/Matrix", "g_h0_lin/bias", "g_bn0/gamma", "d_bn3/beta"} <= f32_names
    assert {"g_h0_lin/Matrix", "g_h1/w", "g_h2/w", "g_h3/w", "d_h1_conv/w", "d_h2_conv/w", "d_h3_conv/w"} == \
        set(specs) - f32_names


def test_named_views_alias_the_flat_buffers():
    cfg = ModelConfig(batch_size=4)
    A = ParamArena(cfg, "cpu")
    views = A.named(A.master)
    assert views["g_h1/w"].shape == (5, 5, 256, 512) and views["d_h1_conv/w"].shape == (5, 5, 64, 128)
    views["d_h3_lin/bias"].fill_(3.0)
    off, _ = A.offsets["d_h3_lin/bias"]
    assert A.master[off].item() == 3.0 and A.master.sum().item() == 3.0
    assert A.bf16.dtype == torch.bfloat16 and A.bf16.numel() == A.master.numel() == A.m.numel() == A.v.numel()


def test_wide_config_arena():
    cfg = ModelConfig(batch_size=2, output_size=128, gf_dim=128, df_dim=128)
    A = ParamArena(cfg, "cpu")
    assert A.named(A.master)["g_h1/w"].shape == (5, 5, 512, 1024)
    assert A.padded % 1024 == 0


def test_optimizer_shards_tile_the_buckets():
    """parallel/fused_comm.py shard_range() mirrors shard_quads() of csrc/comm_sm100.cu: per bucket the ranks own
    consecutive, disjoint ranges that cover it exactly, and every boundary falls on an even number of float4
    quads so that no 16-byte bf16 chunk (8 elements) straddles two owners."""
    from distributed_tensorflow_for_dcgan_b200.parallel.fused_comm import FusedComm
    A = ParamArena(ModelConfig(batch_size=4), "cpu")
    for world in (2, 4, 8):
        comm = FusedComm.__new__(FusedComm)          # only the pure bookkeeping is exercised here
        comm.world = world
        comm.buckets = {"g": (0, A.g_end), "d": (A.g_end, A.padded)}
        for bucket, (lo, hi) in comm.buckets.items():
            cursor = lo
            for r in range(world):
                s0, s1 = comm.shard_range(bucket, r)
                assert s0 == cursor and s1 >= s0, (world, bucket, r)
                assert (s0 - lo) % 8 == 0, (world, bucket, r)
                cursor = s1
            assert cursor == hi, (world, bucket)
            sizes = [comm.shard_range(bucket, r)[1] - comm.shard_range(bucket, r)[0] for r in range(world)]
            assert max(sizes) - min(sizes[:-1] or sizes) <= 8 or sizes[-1] <= max(sizes)   # balanced up to rounding


def test_generator_filters_come_first_and_the_linear_last():
    """The early generator update / exchange covers [g_h1/w, g_h0_lin/Matrix): those three filters must be contiguous at
    the start of the arena and g_h0_lin/Matrix (whose gradient is the last kernel of the step) must follow them,
    before the f32-consumed variables."""
    A = ParamArena(ModelConfig(batch_size=4), "cpu")
    o = {n: A.offsets[n][0] for n in ("g_h1/w", "g_h2/w", "g_h3/w", "g_h0_lin/Matrix")}
    assert o["g_h1/w"] == 0 and o["g_h1/w"] < o["g_h2/w"] < o["g_h3/w"] < o["g_h0_lin/Matrix"] < A.f32_start["g"]
    n3 = 1
    for d in A.offsets["g_h3/w"][1]:
        n3 *= d
    assert o["g_h0_lin/Matrix"] == o["g_h3/w"] + n3                      # nothing between them
    assert o["g_h0_lin/Matrix"] % 8 == 0
    early = o["g_h0_lin/Matrix"] / A.g_end
    assert 0.8 < early < 0.9                                               # ~85 % of the generator's parameters


def test_three_exchange_buckets_tile_the_arena():
    """The generator's bucket is split at g_h0_lin/Matrix; the buckets (g1, g2, d) and their per-rank shards cover
    the padded arena exactly once, and the consumer-side pull range lies inside one bucket."""
    from distributed_tensorflow_for_dcgan_b200.parallel.fused_comm import FusedComm
    A = ParamArena(ModelConfig(batch_size=4), "cpu")
    split = A.offsets["g_h0_lin/Matrix"][0]
    for world in (2, 4, 8):
        comm = FusedComm.__new__(FusedComm)
        comm.world = world
        comm.buckets = {"g1": (0, split), "g2": (split, A.g_end), "d": (A.g_end, A.padded)}
        comm.bucket_names = tuple(comm.buckets)
        cursor = 0
        for b in comm.bucket_names:
            lo, hi = comm.buckets[b]
            assert lo == cursor and (hi - lo) % 4 == 0
            for r in range(world):
                s0, s1 = comm.shard_range(b, r)
                assert s0 == cursor and (s0 - lo) % 8 == 0
                cursor = s1
            assert cursor == hi
        assert cursor == A.padded
        off, shape = A.offsets["g_h0_lin/Matrix"]
        assert comm.buckets["g2"][0] <= off and off + shape[0] * shape[1] <= comm.buckets["g2"][1]


def test_autotune_is_inert_without_a_gpu():
    from distributed_tensorflow_for_dcgan_b200.ops import autotune
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    before = dict(autotune.plans())
    autotune.ensure(autotune.KIND_F, 64, 32, 32, 64, 128)
    assert autotune.plans() == before
