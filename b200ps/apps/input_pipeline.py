"""Image input-pipeline demo — the reference's input_pipeline.py.

CSV ``path,label`` -> first 20 items -> random 5-item test partition -> slice producers
(no shuffle) -> ``read_file`` + ``decode_jpeg(channels=3)`` -> 28x28x3 -> batches of 5; prints 20 train and
10 test label batches (input_pipeline.py:21-29,43-57,60-91,104-110).  The JPEG decoder is the framework's own
(``b200ps.io.jpeg``: Huffman decode on the host, inverse DCT / chroma upsampling / YCbCr->RGB on the GPU, numpy
restatement on a machine without one; bit-exact against libjpeg).  There is no image data offline, so
``make_dataset`` writes small synthetic JPEG files (PIL encoder); ``.raw`` files of 28*28*3 uint8 pixels are
still accepted.
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


def synthetic_image(i: int) -> np.ndarray:
    """A smooth 28x28x3 picture whose look depends on ``i`` (blobs + gradient + a little noise)."""
    rng = np.random.RandomState(i)
    yy, xx = np.mgrid[0:IMAGE_HEIGHT, 0:IMAGE_WIDTH]
    cy, cx = rng.uniform(6, 22, size=2)
    blob = 160.0 * np.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2.0 * (3.0 + i % 5) ** 2))
    img = np.stack([blob + 40, 0.5 * blob + 4.0 * xx, 255 - blob - 3.0 * yy], axis=2) + rng.normal(0, 6, (IMAGE_HEIGHT, IMAGE_WIDTH, 3))
    return np.clip(img, 0, 255).astype(np.uint8)


def make_dataset(directory: str, n: int = 24) -> str:
    """``n`` synthetic JPEG files + the label csv (raw pixel files when no JPEG encoder is importable)."""
    os.makedirs(directory, exist_ok=True)
    try:
        from PIL import Image
    except Exception:                      # pragma: no cover - PIL ships with this image
        Image = None
    with open(os.path.join(directory, "train-labels.csv"), "w") as f:
        for i in range(n):
            img = synthetic_image(i)
            if Image is not None:
                name = f"img_{i:03d}.jpg"
                Image.fromarray(img).save(os.path.join(directory, name), format="JPEG", quality=90,
                                          subsampling=(0, 1, 2)[i % 3])     # 4:4:4, 4:2:2 and 4:2:0 files
            else:
                name = f"img_{i:03d}.raw"
                img.tofile(os.path.join(directory, name))
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


def decode_image(raw: bytes, device: str = "cpu"):
    """``decode_jpeg(channels=3)`` + ``set_shape([28, 28, 3])`` (input_pipeline.py:70-78): a uint8 [28, 28, 3] tensor on
    ``device``.  JPEG streams go through the framework's decoder; anything else is taken as raw pixels."""
    import torch
    if raw[:2] == b"\xff\xd8":
        from ..io.jpeg import decode_jpeg
        img = decode_jpeg(raw, device=device)
    else:
        img = torch.from_numpy(np.frombuffer(raw, dtype=np.uint8).reshape(IMAGE_HEIGHT, IMAGE_WIDTH, NUM_CHANNELS).copy()).to(device)
    if tuple(img.shape) != (IMAGE_HEIGHT, IMAGE_WIDTH, NUM_CHANNELS):
        raise ValueError(f"image is {tuple(img.shape)}, the pipeline is declared for {(IMAGE_HEIGHT, IMAGE_WIDTH, NUM_CHANNELS)}")
    return img


def main(argv=None) -> int:
    root = FLAGS.dataset_path or make_dataset(os.path.join(tempfile.gettempdir(), f"b200ps_jpgs_{os.getuid()}"))
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

    import torch
    device = "cuda" if torch.cuda.is_available() else "cpu"

    def load(rows):
        imgs = torch.stack([decode_image(open(p, "rb").read(), device) for p, _ in rows])
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
