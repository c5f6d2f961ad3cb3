"""Launch script with the reference's name and flags:

    python image_train.py --job_name=ps     --task_index=0 --ps_hosts=h0:2222 --worker_hosts=h1:2222,h2:2222
    python image_train.py --job_name=worker --task_index=0 --ps_hosts=h0:2222 --worker_hosts=h1:2222,h2:2222

(reference: image_train.py:28-35,248-249).  See distributed_tensorflow_for_dcgan_b200/train.py."""
import sys

from distributed_tensorflow_for_dcgan_b200.train import main

if __name__ == "__main__":
    sys.exit(main())
