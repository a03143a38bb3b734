"""Input side of the training step: batches of raw CIFAR-10 records as the kernels want them --
uint8 [B,3,32,32] CHW payloads plus int32 labels (cifar10cnn.py:54-62).  The crop / transpose / cast
of read_cifar_files (:65-68) is done inside the first CUDA kernel, not here.

Two sources: the binary files of cifar10cnn.py:78-80 when they are already on disk (there is no
network: download_data :34-52 is out of scope), else synthetic records of the same shape."""
from __future__ import annotations

import os

import numpy as np
import torch

RECORD_LENGTH = 3073          # cifar10cnn.py:24-25: 1 label byte + 32*32*3 image bytes
DATA_DIR = "cifar10data"      # :26 (the --data_dir flag is parsed but unused by the reference)
EXTRACT_FOLDER = "cifar-10-batches-bin"


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


class BatchSource:
    """shuffle_batch stand-in (:87-90): every worker samples its own stream independently (the
    reference does not shard the epoch between workers).  Records live in HBM (150 MB for
    the whole training set); a batch is an index gather on the device."""

    def __init__(self, batch_size, train_logical=True, device="cuda", seed=0, data_dir=DATA_DIR, synthetic=None,
                 synthetic_records=8192):
        self.batch_size = batch_size
        files = cifar_files(train_logical, data_dir)
        have = all(os.path.isfile(f) for f in files)
        self.synthetic = (not have) if synthetic is None else synthetic
        g = np.random.default_rng(seed)
        if self.synthetic:
            images = g.integers(0, 256, (synthetic_records, 3, 32, 32), dtype=np.uint8)
            labels = g.integers(0, 10, (synthetic_records,)).astype(np.int32)
        else:
            images, labels = read_records(files)
        self.images = torch.as_tensor(images).to(device)
        self.labels = torch.as_tensor(labels).to(device)
        self.gen = torch.Generator(device=device)
        self.gen.manual_seed(seed)

    def next(self):
        idx = torch.randint(0, self.images.shape[0], (self.batch_size,), device=self.images.device, generator=self.gen)
        return self.images[idx].contiguous(), self.labels[idx].contiguous()
