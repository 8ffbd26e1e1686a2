"""Image input-pipeline demo — the reference's input_pipeline.py, minus the JPEG codec.

CSV ``path,label`` -> first 20 items -> random 5-item test partition -> slice producers
(no shuffle) -> file read -> decode to 28x28x3 -> batches of 5; prints 20 train and 10 test
label batches (input_pipeline.py:21-29,43-57,60-91,104-110).  There is no image data (or
libjpeg binding) in this environment, so files hold raw 28*28*3 uint8 pixels (``.raw``)
written by ``make_dataset``; a real decoder plugs into ``decode_image``.
"""
from __future__ import annotations

import os
import random
import tempfile

import numpy as np

from .. import flags
from ..data.queue import Batcher, Coordinator, slice_input_producer, start_queue_runners

FLAGS = flags.FLAGS
flags.DEFINE_string("dataset_path", "", "directory with image files + the label csv (made on the fly when empty)")
flags.DEFINE_string("label_file", "train-labels.csv", "csv of 'path,label' lines")

IMAGE_HEIGHT, IMAGE_WIDTH, NUM_CHANNELS, BATCH_SIZE, TEST_SET_SIZE = 28, 28, 3, 5, 5


def make_dataset(directory: str, n: int = 24) -> str:
    os.makedirs(directory, exist_ok=True)
    rng = np.random.RandomState(0)
    with open(os.path.join(directory, "train-labels.csv"), "w") as f:
        for i in range(n):
            name = f"img_{i:03d}.raw"
            rng.randint(0, 256, size=IMAGE_HEIGHT * IMAGE_WIDTH * NUM_CHANNELS, dtype=np.uint8).tofile(os.path.join(directory, name))
            f.write(f"{name},{i % 10}\n")
    return directory


def encode_label(label):
    return int(label)


def read_label_file(path):
    filepaths, labels = [], []
    with open(path) as f:
        for line in f:
            fp, lab = line.strip().split(",")
            filepaths.append(fp)
            labels.append(encode_label(lab))
    return filepaths, labels


def decode_image(raw: bytes) -> np.ndarray:
    return np.frombuffer(raw, dtype=np.uint8).reshape(IMAGE_HEIGHT, IMAGE_WIDTH, NUM_CHANNELS)


def main(argv=None) -> int:
    root = FLAGS.dataset_path or make_dataset(os.path.join(tempfile.gettempdir(), f"b200ps_imgs_{os.getuid()}"))
    filepaths, labels = read_label_file(os.path.join(root, FLAGS.label_file))
    filepaths = [os.path.join(root, fp) for fp in filepaths][:20]
    labels = labels[:20]
    # random partition (dynamic_partition with a shuffled 0/1 vector, input_pipeline.py:51-57)
    partitions = [0] * len(filepaths)
    partitions[:TEST_SET_SIZE] = [1] * TEST_SET_SIZE
    random.shuffle(partitions)
    train = [(f, l) for f, l, p in zip(filepaths, labels, partitions) if p == 0]
    test = [(f, l) for f, l, p in zip(filepaths, labels, partitions) if p == 1]
    tq = slice_input_producer(list(zip(*train)), shuffle=False, name="train_slices")
    vq = slice_input_producer(list(zip(*test)), shuffle=False, name="test_slices")
    train_batch = Batcher(tq, BATCH_SIZE, name="train_batch")
    test_batch = Batcher(vq, BATCH_SIZE, name="test_batch")
    print("input pipeline ready")
    coord = Coordinator()
    threads = start_queue_runners(coord)

    def load(rows):
        imgs = np.stack([decode_image(open(p, "rb").read()) for p, _ in rows])
        return imgs, [l for _, l in rows]

    print("from the train set:")
    for _ in range(20):
        imgs, labs = load(train_batch.next(timeout=4.0))
        print(labs)
    print("from the test set:")
    for _ in range(10):
        imgs, labs = load(test_batch.next(timeout=4.0))
        print(labs)
    for q in (tq, vq, train_batch.queue, test_batch.queue):
        q.close(cancel_pending_enqueues=True)
    coord.request_stop()
    coord.join(threads)
    return 0


if __name__ == "__main__":
    flags.run(main)
