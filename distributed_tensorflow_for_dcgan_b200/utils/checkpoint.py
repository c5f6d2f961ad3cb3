"""Checkpoint save / restore keyed by the reference's ``tf.train.Saver`` variable names.

The reference checkpoints through ``tf.train.Supervisor(saver=tf.train.Saver(), logdir=
checkpoint_dir, save_model_secs=600)`` and restores with ``get_checkpoint_state`` +
``saver.restore`` (image_train.py:103,123-129,233-245).  On disk that is ``model.ckpt-<step>``
plus a text-proto ``checkpoint`` state file.  Binary compatibility with the TF-0.x tensor-slice
format is not a goal; the contract is the *name -> tensor* mapping in TF layouts:

* weights: ``g_h1/w`` = ``[5,5,Cout,Cin]``, ``d_h1_conv/w`` = ``[5,5,Cin,Cout]``, ``*/Matrix`` =
  ``[in,out]`` ... exactly the manifest of SURVEY.md 2.5;
* ``Variable`` = global_step (int32);  EMA shadows under the canonical names of
  ``models/manifest.py``; Adam slots under TF slot names (extra keys, see manifest).

Tensors are stored with safetensors (``model.ckpt-<step>.safetensors``); the ``checkpoint`` state
file uses TensorFlow's text format so tooling that only looks for the newest path keeps working.
"""
from __future__ import annotations

import os
import re
from typing import Dict, List, Optional

import torch

STATE_FILE = "checkpoint"
PREFIX = "model.ckpt"
SUFFIX = ".safetensors"


class CheckpointState:
    def __init__(self, model_checkpoint_path: str, all_model_checkpoint_paths: List[str]):
        self.model_checkpoint_path = model_checkpoint_path
        self.all_model_checkpoint_paths = all_model_checkpoint_paths


def _write_state(ckpt_dir: str, paths: List[str]) -> None:
    tmp = os.path.join(ckpt_dir, STATE_FILE + ".tmp")
    with open(tmp, "w") as f:
        f.write(f'model_checkpoint_path: "{paths[-1]}"\n')
        for p in paths:
            f.write(f'all_model_checkpoint_paths: "{p}"\n')
    os.replace(tmp, os.path.join(ckpt_dir, STATE_FILE))


def get_checkpoint_state(ckpt_dir: str) -> Optional[CheckpointState]:
    """``tf.train.get_checkpoint_state`` (image_train.py:239): None when there is no state file."""
    path = os.path.join(ckpt_dir, STATE_FILE)
    if not os.path.exists(path):
        return None
    latest, allp = None, []
    with open(path) as f:
        for line in f:
            m = re.match(r'\s*(\w+):\s*"(.*)"\s*$', line)
            if not m:
                continue
            if m.group(1) == "model_checkpoint_path":
                latest = m.group(2)
            elif m.group(1) == "all_model_checkpoint_paths":
                allp.append(m.group(2))
    if latest is None:
        return None

    def absolutise(p):
        return p if os.path.isabs(p) else os.path.join(ckpt_dir, p)
    return CheckpointState(absolutise(latest), [absolutise(p) for p in allp])


def save(ckpt_dir: str, state: Dict[str, torch.Tensor], global_step: int, max_to_keep: int = 5) -> str:
    """Writes ``model.ckpt-<global_step>`` and updates the state file; keeps the newest 5 (TF default)."""
    from safetensors.torch import save_file
    os.makedirs(ckpt_dir, exist_ok=True)
    stem = f"{PREFIX}-{int(global_step)}"
    path = os.path.join(ckpt_dir, stem + SUFFIX)
    tensors = {k: v.detach().to("cpu").contiguous() for k, v in state.items()}
    tmp = path + ".tmp"
    save_file(tensors, tmp)
    os.replace(tmp, path)
    st = get_checkpoint_state(ckpt_dir)
    rel = [os.path.basename(p) for p in (st.all_model_checkpoint_paths if st else [])]
    rel = [p for p in rel if p != stem] + [stem]
    while len(rel) > max_to_keep:
        old = rel.pop(0)
        try:
            os.remove(os.path.join(ckpt_dir, old + SUFFIX))
        except OSError:
            pass
    _write_state(ckpt_dir, rel)
    return os.path.join(ckpt_dir, stem)


def restore(model_checkpoint_path: str, device="cpu") -> Dict[str, torch.Tensor]:
    from safetensors.torch import load_file
    path = model_checkpoint_path if model_checkpoint_path.endswith(SUFFIX) else model_checkpoint_path + SUFFIX
    return load_file(path, device=str(device))


def load_latest(ckpt_dir: str, device="cpu") -> Optional[Dict[str, torch.Tensor]]:
    """The reference's ``load()`` (image_train.py:233-245): newest checkpoint of the directory or None."""
    print(" [*] Reading checkpoints...")
    st = get_checkpoint_state(ckpt_dir)
    if st and st.model_checkpoint_path:
        return restore(st.model_checkpoint_path, device)
    return None
