"""Input side of the training step: batches of raw CIFAR-10 records as the kernels want them --
uint8 [B,3,32,32] CHW payloads plus int32 labels (cifar10cnn.py:54-62).  The crop / transpose / cast
of read_cifar_files (:65-68) is done inside the first CUDA kernel, not here.

Two sources: the binary files of cifar10cnn.py:78-80 when they are already on disk (there is no
network: download_data :34-52 is out of scope), else synthetic records of the same shape.

The record set lives in HBM (150 MB for the whole training set) and `tf.train.shuffle_batch` (:87-90:
capacity 5384, min_after_dequeue 5000, one reader thread over a per-epoch shuffled file list :82) is one
CUDA kernel per batch (csrc/shuffle_batch.cu through b200_shuffle_batch): a uniform draw of `batch` of the
5384 queued record ids without replacement, the gather of those records, and the reader's refill with the next
sequential records.  Every worker runs its own queue over all files with its own seed -- the reference does not
shard the epoch between workers.
"""
from __future__ import annotations

import os

import numpy as np
import torch

RECORD_LENGTH = 3073          # cifar10cnn.py:24-25: 1 label byte + 32*32*3 image bytes
DATA_DIR = "cifar10data"      # :26 (the --data_dir flag is parsed but unused by the reference)
EXTRACT_FOLDER = "cifar-10-batches-bin"
QUEUE_CAPACITY = 5384         # :84-85: min_after_dequeue 5000 + 3 * batch_size
MIN_AFTER_DEQUEUE = 5000


def cifar_files(train_logical=True, data_dir=DATA_DIR):
    """file list of input_pipeline (:77-80)."""
    if train_logical:
        return [os.path.join(data_dir, EXTRACT_FOLDER, f"data_batch_{i}.bin") for i in range(1, 6)]
    return [os.path.join(data_dir, EXTRACT_FOLDER, "test_batch.bin")]


def read_records(files):
    """FixedLengthRecordReader(3073) + decode_raw + slice (:55-62) for whole files at once."""
    raw = np.concatenate([np.fromfile(f, dtype=np.uint8) for f in files])
    if raw.size % RECORD_LENGTH:
        raise ValueError("CIFAR binary file size is not a multiple of 3073 bytes")
    rec = raw.reshape(-1, RECORD_LENGTH)
    return rec[:, 1:].reshape(-1, 3, 32, 32).copy(), rec[:, 0].astype(np.int32)


def epoch_file_orders(n_files, n_epochs, seed):
    """string_input_producer(files) (:82, shuffle=True by default): a fresh random order of the file list per epoch,
    flattened [n_epochs * n_files]."""
    rng = np.random.default_rng(seed)
    return np.concatenate([rng.permutation(n_files) for _ in range(n_epochs)]).astype(np.int32)


class BatchSource:
    """shuffle_batch (:87-90) on the device.  next() returns the views of one of two alternating output buffers
    (valid until the next-but-one call); next_into() writes into caller buffers."""

    def __init__(self, batch_size, train_logical=True, device="cuda", seed=0, data_dir=DATA_DIR, synthetic=None,
                 synthetic_records=8192, capacity=None, n_epochs=256):
        from . import _lib
        self.lib = _lib.load()
        self.batch_size = batch_size
        files = cifar_files(train_logical, data_dir)
        have = all(os.path.isfile(f) for f in files)
        self.synthetic = (not have) if synthetic is None else synthetic
        g = np.random.default_rng(seed)
        if self.synthetic:
            images = g.integers(0, 256, (synthetic_records, 3, 32, 32), dtype=np.uint8)
            labels = g.integers(0, 10, (synthetic_records,)).astype(np.int32)
            n_files = 1
        else:
            sizes = {os.path.getsize(f) for f in files}
            if len(sizes) != 1:
                raise ValueError("the CIFAR binary files must all have the same number of records")
            images, labels = read_records(files)
            n_files = len(files)
        self.n_records = int(images.shape[0])
        self.per_file = self.n_records // n_files
        # capacity = min_after_dequeue + 3 * batch (:84-85); shrunk only for record sets smaller than the queue
        cap = MIN_AFTER_DEQUEUE + 3 * batch_size if capacity is None else capacity
        self.capacity = int(min(cap, self.n_records, 12000))
        if self.capacity <= batch_size:
            raise ValueError("record set too small for the shuffle queue")
        self.seed = int(seed)
        self.file_order_host = epoch_file_orders(n_files, n_epochs, seed)
        self.images = torch.as_tensor(images).reshape(self.n_records, -1).contiguous().to(device)
        self.labels = torch.as_tensor(labels).to(device)
        self.file_order = torch.as_tensor(self.file_order_host).to(device)
        pos = np.arange(self.capacity, dtype=np.int64)
        first = self.file_order_host[(pos // self.per_file) % self.file_order_host.size].astype(np.int64) * self.per_file \
            + pos % self.per_file
        self.slots = [torch.as_tensor(first.astype(np.int32)).to(device), torch.empty(self.capacity, dtype=torch.int32,
                                                                                       device=device)]
        self.step = 0
        self.stream_pos = self.capacity
        self.image_bytes = int(self.images.shape[1])
        self._out = [(torch.empty(batch_size, 3, 32, 32, dtype=torch.uint8, device=device),
                      torch.empty(batch_size, dtype=torch.int32, device=device)) for _ in range(2)]

    def next_into(self, images_out, labels_out, stream=None):
        """one dequeue_many(batch) + refill, enqueued on `stream` (default: the current stream)."""
        from ._lib import check
        st = torch.cuda.current_stream().cuda_stream if stream is None else stream
        k = self.step & 1
        check(self.lib.b200_shuffle_batch(self.images.data_ptr(), self.labels.data_ptr(), self.n_records,
                                          self.file_order.data_ptr(), int(self.file_order.numel()), self.per_file,
                                          self.slots[k].data_ptr(), self.slots[1 - k].data_ptr(), self.capacity,
                                          self.seed, self.step, self.stream_pos, images_out.data_ptr(),
                                          labels_out.data_ptr(), self.batch_size, self.image_bytes, st))
        self.step += 1
        self.stream_pos += self.batch_size

    def next(self):
        img, lab = self._out[self.step & 1]
        self.next_into(img, lab)
        return img, lab
