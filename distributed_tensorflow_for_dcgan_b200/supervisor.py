"""Supervision: chief election, init-or-restore, wait-for-chief, timed checkpoints and summaries,
stop conditions.

Stands in for ``tf.train.Supervisor(is_chief=task_index==0, logdir=checkpoint_dir, saver=saver,
global_step=global_step, save_model_secs=600)`` + ``sv.managed_session`` + ``sv.should_stop`` +
``sv.summary_computed`` and the explicit ``load()`` (image_train.py:123-129,141-146,150,174,233-245).

Differences, on purpose: checkpoints are taken from the training loop at a step boundary (the
Supervisor's timer thread snapshots while variables are being updated); the Adam slots are saved;
SIGINT/SIGTERM request a clean stop, and the chief writes a final checkpoint on the way out.
"""
from __future__ import annotations

import signal
import threading
import time
from typing import Optional

from .utils import checkpoint as ckpt
from .utils.summary import SummaryWriter


class Supervisor:
    def __init__(self, is_chief: bool, logdir: str, save_model_secs: int = 600, save_summaries_secs: int = 10,
                 max_steps: int = 1200000, install_signal_handlers: bool = True):
        self.is_chief, self.logdir = is_chief, logdir
        self.save_model_secs, self.save_summaries_secs, self.max_steps = save_model_secs, save_summaries_secs, max_steps
        self._stop = threading.Event()
        self._writer: Optional[SummaryWriter] = None
        self.start_time = time.time()
        self._next_ckpt = self.start_time + save_model_secs
        self._next_summary = self.start_time + save_summaries_secs
        self.last_checkpoint_path: Optional[str] = None
        if install_signal_handlers and threading.current_thread() is threading.main_thread():
            for sig in (signal.SIGINT, signal.SIGTERM):
                try:
                    signal.signal(sig, lambda *_: self.request_stop())
                except (ValueError, OSError):
                    pass

    # ---- session bring-up (image_train.py:141-146) -----------------------------------------------------------
    def init_or_restore(self, trainer) -> bool:
        """Chief: restore the newest checkpoint of ``logdir`` if there is one (the Supervisor's own restore
        plus the explicit ``load()``), else keep the fresh initialisation.  Returns True when restored."""
        if not self.is_chief:
            return False
        state = ckpt.load_latest(self.logdir)
        if state is not None:
            trainer.load_state_dict(state)
            print("load success!")
            return True
        print("load failed!!")
        return False

    # ---- stop conditions (image_train.py:150) -------------------------------------------------------------------
    def should_stop(self) -> bool:
        return self._stop.is_set()

    def request_stop(self) -> None:
        self._stop.set()

    def keep_going(self, step: int) -> bool:
        """``not sv.should_stop() and step < 1200000``."""
        return not self.should_stop() and step < self.max_steps

    # ---- timed services --------------------------------------------------------------------------------------------
    def summary_due(self) -> bool:
        """``next_summary_time > time.time()`` inverted (image_train.py:155)."""
        return time.time() >= self._next_summary

    def summary_computed(self, summary: bytes, global_step: int) -> None:
        if self.is_chief:
            print("Running Summary operation on the chief.")
            if self._writer is None:
                self._writer = SummaryWriter(self.logdir)
            self._writer.add_summary(summary, global_step)
            print("Finished running Summary operation.")
        self._next_summary += self.save_summaries_secs  # image_train.py:178

    def maybe_checkpoint(self, trainer, global_step: int, force: bool = False) -> Optional[str]:
        if not self.is_chief:
            return None
        now = time.time()
        if not force and now < self._next_ckpt:
            return None
        self._next_ckpt = now + self.save_model_secs
        self.last_checkpoint_path = ckpt.save(self.logdir, trainer.state_dict(), global_step)
        return self.last_checkpoint_path

    def close(self) -> None:
        if self._writer is not None:
            self._writer.close()
