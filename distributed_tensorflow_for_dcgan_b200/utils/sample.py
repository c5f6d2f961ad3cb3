"""Sample grid writer: ``save_images`` / ``merge`` / ``inverse_transform`` (image_train.py:197-219).

``scipy.misc.imsave`` (gone from modern SciPy) rescaled the array's min..max to 0..255 before
encoding; that behaviour is kept.  The reference's debugging ``print len(image)`` per tile (:210)
is dropped.  On the B200 path the tiling runs on the device (``grid_merge`` kernel) and only the
finished grid crosses PCIe.
"""
from __future__ import annotations

import os

import numpy as np


def inverse_transform(images: np.ndarray) -> np.ndarray:
    return (images + 1.0) / 2.0


def merge(images: np.ndarray, size) -> np.ndarray:
    """Tiles N images row-major into one ``[rows * h, cols * w, c]`` grid (``size = (rows, cols)``); cells beyond N stay
    black.  One reshape / transpose instead of a per-image copy loop."""
    rows, cols = int(size[0]), int(size[1])
    n, h, w, c = images.shape
    cells = np.zeros((rows * cols, h, w, c), dtype=np.float64)
    cells[:min(n, rows * cols)] = images[:rows * cols]
    return cells.reshape(rows, cols, h, w, c).transpose(0, 2, 1, 3, 4).reshape(rows * h, cols * w, c)


def to_uint8_minmax(arr: np.ndarray) -> np.ndarray:
    lo, hi = float(arr.min()), float(arr.max())
    scale = 255.0 / (hi - lo) if hi > lo else 1.0
    return ((arr - lo) * scale).round().clip(0, 255).astype(np.uint8)


def imsave(path: str, arr: np.ndarray) -> None:
    from PIL import Image
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    Image.fromarray(to_uint8_minmax(arr)).save(path)


def save_images(images: np.ndarray, size, image_path: str) -> None:
    imsave(image_path, merge(inverse_transform(images), size))


def save_merged(grid: np.ndarray, image_path: str) -> None:
    """For a grid that was already inverse-transformed and tiled on the device."""
    imsave(image_path, grid)


def sample_path(sample_dir: str, step: int, steps_per_epoch: int) -> str:
    """``'./samples/train_{:02d}_{:04d}.png'.format(step / step_per_epoch, step % step_per_epoch)``
    (image_train.py:188-190, Python-2 integer division); the directory follows ``--sample_dir``."""
    return os.path.join(sample_dir, "train_{:02d}_{:04d}.png".format(step // steps_per_epoch, step % steps_per_epoch))
