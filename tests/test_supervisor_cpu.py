"""Supervisor semantics: chief-only services, init-or-restore, time-gated checkpoints and
summaries, stop conditions -- the stand-in for tf.train.Supervisor in image_train.py:123-129,141-150,174."""
import os
import signal
import time

import torch

from distributed_tensorflow_for_dcgan_b200.supervisor import Supervisor
from distributed_tensorflow_for_dcgan_b200.utils import checkpoint
from distributed_tensorflow_for_dcgan_b200.utils.summary import Summary, read_events


class _Trainer:
    """Anything with state_dict / load_state_dict."""

    def __init__(self):
        self.state = {"w": torch.arange(4.0), "Variable": torch.tensor(0, dtype=torch.int32)}

    def state_dict(self):
        return {k: v.clone() for k, v in self.state.items()}

    def load_state_dict(self, sd):
        self.state = {k: v.clone() for k, v in sd.items()}


def test_stop_conditions(tmp_path):
    sv = Supervisor(True, str(tmp_path), max_steps=10, install_signal_handlers=False)
    assert sv.keep_going(0) and sv.keep_going(9)
    assert not sv.keep_going(10)                      # step < max_steps (image_train.py:150)
    sv.request_stop()
    assert sv.should_stop() and not sv.keep_going(0)


def test_sigterm_requests_a_clean_stop(tmp_path):
    old = signal.getsignal(signal.SIGTERM)
    try:
        sv = Supervisor(True, str(tmp_path), install_signal_handlers=True)
        os.kill(os.getpid(), signal.SIGTERM)
        t0 = time.time()
        while not sv.should_stop() and time.time() - t0 < 5:
            time.sleep(0.01)
        assert sv.should_stop()
    finally:
        signal.signal(signal.SIGTERM, old)
        signal.signal(signal.SIGINT, signal.default_int_handler)


def test_checkpoint_timer_chief_only_and_restore(tmp_path):
    tr = _Trainer()
    chief = Supervisor(True, str(tmp_path), save_model_secs=3600, install_signal_handlers=False)
    other = Supervisor(False, str(tmp_path), save_model_secs=0, install_signal_handlers=False)
    assert chief.maybe_checkpoint(tr, 1) is None      # not due yet
    assert other.maybe_checkpoint(tr, 1, force=True) is None   # non-chief never writes
    assert not other.init_or_restore(tr)
    assert not chief.init_or_restore(tr)              # nothing to restore: "load failed!!", fresh init kept
    tr.state["w"] += 1
    tr.state["Variable"] = torch.tensor(7, dtype=torch.int32)
    path = chief.maybe_checkpoint(tr, 7, force=True)
    assert path and os.path.basename(path).startswith("model.ckpt-7")
    st = checkpoint.get_checkpoint_state(str(tmp_path))
    assert st is not None and st.model_checkpoint_path.endswith("model.ckpt-7")
    # a restarted chief resumes from it
    tr2 = _Trainer()
    sv2 = Supervisor(True, str(tmp_path), install_signal_handlers=False)
    assert sv2.init_or_restore(tr2)
    assert torch.equal(tr2.state["w"], tr.state["w"]) and int(tr2.state["Variable"]) == 7
    # due timer: save_model_secs = 0 -> every call writes
    eager = Supervisor(True, str(tmp_path), save_model_secs=0, install_signal_handlers=False)
    assert eager.maybe_checkpoint(tr, 8) is not None


def test_summary_timer_and_event_file(tmp_path, capsys):
    sv = Supervisor(True, str(tmp_path), save_summaries_secs=3600, install_signal_handlers=False)
    assert not sv.summary_due()
    sv._next_summary = time.time() - 1                # the tick has passed
    assert sv.summary_due()
    s = Summary()
    s.scalar("d_loss", 1.5)
    sv.summary_computed(s.serialize(), 3)
    assert not sv.summary_due()                       # next tick is save_summaries_secs later (image_train.py:178)
    out = capsys.readouterr().out
    assert "Running Summary operation on the chief." in out and "Finished running Summary operation." in out
    sv.close()
    files = [f for f in os.listdir(tmp_path) if "tfevents" in f]
    assert len(files) == 1
    events = list(read_events(os.path.join(tmp_path, files[0])))
    assert any(step == 3 and tags.get("d_loss", (None, None))[0] == "scalar" for step, tags in events)
    # a non-chief computes the summary on the tick and discards it
    nc = Supervisor(False, str(tmp_path / "nc"), save_summaries_secs=1, install_signal_handlers=False)
    nc.summary_computed(s.serialize(), 3)
    assert not os.path.exists(tmp_path / "nc")
