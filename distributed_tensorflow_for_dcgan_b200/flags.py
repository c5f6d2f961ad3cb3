"""Command-line flags: every flag name of the reference (image_train.py:10-40) with its default,
plus the new ones of SURVEY.md 5.6.  Of the flags the reference declares but never reads:

* live here: ``batch_size``, ``output_size``, ``c_dim`` (model / batch geometry), ``sample_dir`` (where the grids
  go), ``train_size`` (a finite value shortens the epoch used by the log line and the sample file names,
  ``config.steps_per_epoch``) and ``log_device_placement`` (the chief prints which rank / device holds every
  variable, ``train.log_placement``);
* accepted and unused, as in the reference: ``epoch`` (the loop budget is ``--max_steps`` = the reference's
  hard-coded 1,200,000 global steps), ``image_size``, ``dataset``, ``is_train``, ``is_crop``, ``visualize``.
"""
from __future__ import annotations

import math
import pprint

from absl import flags

FLAGS = flags.FLAGS
_DEFINED = False


def define_flags() -> None:
    global _DEFINED
    if _DEFINED:
        return
    _DEFINED = True
    # ---- reference flags (image_train.py:11-38) ---------------------------------------------------------
    flags.DEFINE_integer("epoch", 25, "Epoch to train [25]")
    flags.DEFINE_float("learning_rate", 0.0002, "Learning rate of for adam [0.0002]")
    flags.DEFINE_float("beta1", 0.5, "Momentum term of adam [0.5]")
    flags.DEFINE_float("train_size", math.inf, "The size of train images [np.inf]")
    flags.DEFINE_integer("batch_size", 64, "The size of batch images [64]")
    flags.DEFINE_integer("image_size", 108, "The size of image to use (will be center cropped) [108]")
    flags.DEFINE_integer("output_size", 64, "The size of the output images to produce [64]")
    flags.DEFINE_integer("c_dim", 3, "Dimension of image color. [3]")
    flags.DEFINE_string("dataset", "celebA", "The name of dataset [celebA, mnist, lsun]")
    flags.DEFINE_string("checkpoint_dir", "checkpoint", "Directory name to save the checkpoints [checkpoint]")
    flags.DEFINE_string("sample_dir", "samples", "Directory name to save the image samples [samples]")
    flags.DEFINE_boolean("is_train", False, "True for training, False for testing [False]")
    flags.DEFINE_boolean("is_crop", False, "True for training, False for testing [False]")
    flags.DEFINE_boolean("visualize", False, "True for visualizing, False for nothing [False]")
    flags.DEFINE_string("data_dir", "train", "dir for train set")
    flags.DEFINE_string("sample_image_dir", "sample_data", "dir for train set")
    flags.DEFINE_string("ps_hosts", "", "Comma-separated list of hostname:port pairs")
    flags.DEFINE_string("worker_hosts", "", "Comma-separated list of hostname:port pairs")
    flags.DEFINE_string("job_name", "", "One of 'ps', 'worker'")
    flags.DEFINE_integer("task_index", 0, "Index of task within the job")
    flags.DEFINE_bool("log_device_placement", True, "whether to log the device placement")
    flags.DEFINE_integer("save_summaries_secs", 10, "Save summaries interval seconds.")
    # ---- new flags -------------------------------------------------------------------------
    flags.DEFINE_bool("sync_replicas", True, "synchronous replicas (average gradients, one Adam apply per "
                      "step); False = asynchronous parameter-server mode like the reference")
    flags.DEFINE_enum("update_mode", "simultaneous", ["simultaneous", "alternating"],
                      "D and G from one forward pass (reference) or D then G")
    flags.DEFINE_enum("device", "auto", ["auto", "cuda", "cpu"], "cuda = B200 engine, cpu = PyTorch/gloo plumbing")
    flags.DEFINE_enum("dtype", "bf16", ["bf16", "fp32"], "compute dtype of the CUDA engine (cpu path is fp32)")
    flags.DEFINE_integer("max_steps", 1200000, "global-step budget (image_train.py:150)")
    flags.DEFINE_bool("synthetic_data", False, "synthetic images instead of TFRecords")
    flags.DEFINE_integer("ngf", 64, "generator base width (gf_dim)")
    flags.DEFINE_integer("ndf", 64, "discriminator base width (df_dim)")
    flags.DEFINE_integer("z_dim", 100, "latent dimension")
    flags.DEFINE_bool("sync_bn", False, "cross-replica batch-norm statistics")
    flags.DEFINE_enum("comm_backend", "fused", ["fused", "nccl"], "multi-GPU gradient path")
    flags.DEFINE_integer("save_model_secs", 600, "checkpoint period (image_train.py:129)")
    flags.DEFINE_integer("sample_every", 100, "sample-grid period in global steps (image_train.py:179)")
    flags.DEFINE_integer("seed", 0, "initialisation / data seed")
    flags.DEFINE_bool("cuda_graph", True, "replay the step from a CUDA graph")
    flags.DEFINE_integer("num_reader_threads", 16, "input pipeline threads (image_input.py:76)")
    flags.DEFINE_bool("gpu_decode", True, "GPU runs: copy the float64 records to the device as stored and cast them "
                      "there (image_input.py:48-51 as device work) instead of casting on the reader threads")
    flags.DEFINE_integer("shuffle_min_after_dequeue", -1, "override of int(107766*0.1); -1 keeps the reference value")
    flags.DEFINE_bool("profile", False, "NVTX ranges + CUDA-event step timing report")
    flags.DEFINE_bool("identical_replicas", False, "testing aid: every worker draws the same data and z, so "
                      "synchronous replicas must reproduce the single-process run exactly")
    flags.DEFINE_integer("worker_timeout_secs", 120, "ps role: a worker that stays silent this long while others wait "
                         "for it (sync round / shutdown) is presumed dead and dropped; 0 = wait forever like "
                         "server.join() (image_train.py:62-63)")
    flags.DEFINE_integer("peer_timeout_secs", 600, "GPU replicas: how long a rank waits for its peers' signals inside "
                         "the fused exchange before it raises the device error flag and the job stops (the chief "
                         "pauses between steps for checkpoints and summaries)")
    flags.DEFINE_integer("log_every", 1, "print the step line every N steps (reference: every step)")


UNUSED_IN_REFERENCE_AND_HERE = ("epoch", "image_size", "dataset", "is_train", "is_crop", "visualize")


def pprint_flags() -> None:
    """``pp.pprint(flags.FLAGS.__flags)`` (image_train.py:223)."""
    own = {}
    for module, fl in FLAGS.flags_by_module_dict().items():
        if "dcgan" in module or module.endswith("flags"):
            own.update({f.name: f.value for f in fl})
    pprint.pprint(own)


def model_config_from_flags():
    from .config import ModelConfig
    return ModelConfig(output_size=FLAGS.output_size, gf_dim=FLAGS.ngf, df_dim=FLAGS.ndf, c_dim=FLAGS.c_dim,
                       z_dim=FLAGS.z_dim, batch_size=FLAGS.batch_size)
