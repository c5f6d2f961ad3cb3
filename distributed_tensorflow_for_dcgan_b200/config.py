"""Model / data hyper-parameters.

The reference hard-codes these as module constants (``distriubted_model.py:7-11``:
``output_size=64, gf_dim=64, df_dim=64, batch_size=64, c_dim=3``; ``image_train.py:42-48``:
``z_dim=100, sample_size=64, image_num=107766*3``) and leaves the like-named flags dead
(SURVEY.md 1 / 5.6).  Here they are one dataclass that the flags populate, so the dead flags
become live while the defaults reproduce the reference exactly.
"""
from __future__ import annotations

import dataclasses


@dataclasses.dataclass(frozen=True)
class ModelConfig:
    output_size: int = 64     # distriubted_model.py:7
    gf_dim: int = 64          # distriubted_model.py:8
    df_dim: int = 64          # distriubted_model.py:9
    c_dim: int = 3            # distriubted_model.py:11
    z_dim: int = 100          # image_train.py:42
    batch_size: int = 64      # distriubted_model.py:10 / image_train.py:45 (per worker)
    bn_eps: float = 1e-5      # distriubted_model.py:18
    bn_momentum: float = 0.9  # distriubted_model.py:18 (EMA decay)
    lrelu_leak: float = 0.2   # distriubted_model.py:156

    @property
    def s16(self) -> int:
        return self.output_size // 16

    @property
    def g_h0_features(self) -> int:
        return self.gf_dim * 8 * self.s16 * self.s16

    @property
    def d_flat_features(self) -> int:
        return self.df_dim * 8 * self.s16 * self.s16


# image_train.py:43-48 -- epoch arithmetic used only for the log line and sample file names.
IMAGE_NUM = 107766 * 3
NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN = 107766  # image_input.py:14


def steps_per_epoch(batch_size: int = 64, train_size: float = float("inf")) -> int:
    """``step_per_epoch = image_num / batch_size`` under Python-2 integer division = 5051.  A finite
    ``--train_size`` (declared but never read by the reference, image_train.py:14) caps the image count."""
    n = IMAGE_NUM if train_size == float("inf") else max(int(min(IMAGE_NUM, train_size)), batch_size)
    return n // batch_size


WIDE_128 = ModelConfig(output_size=128, gf_dim=128, df_dim=128)
