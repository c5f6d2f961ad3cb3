# End-to-end on one GPU with real TFRecord files: writes a small float64 dataset, trains a few steps through
# image_train.py with the device-side decode (default) and with the reader-thread cast, prints the last log lines.
set -e
D=$(mktemp -d)
python - <<PY
import numpy as np, os
from distributed_tensorflow_for_dcgan_b200.data import tfrecord
rng = np.random.RandomState(0)
os.makedirs("$D/train")
for i in range(2):
    tfrecord.write_image_tfrecord("$D/train/image_%d.tfrecords" % i, rng.rand(160, 64, 64, 3) * 2 - 1)
PY
for dec in true false; do
  echo "== gpu_decode=$dec"
  python image_train.py --data_dir=$D/train --sample_image_dir=$D/train --checkpoint_dir=$D/ckpt_$dec --sample_dir=$D/samples_$dec \
      --shuffle_min_after_dequeue=64 --max_steps=12 --gpu_decode=$dec --sample_every=10 --num_reader_threads=4 2>&1 | tail -4
done
