from .adam_tf import AdamTF, adam_tf_step_  # noqa: F401
