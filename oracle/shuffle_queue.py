"""Host model (TEST INFRASTRUCTURE ONLY) of the input queue of cifar10cnn.py:72-91.

`tf.train.shuffle_batch(capacity=5384, min_after_dequeue=5000, num_threads=1)` (:87-90) over
`tf.train.string_input_producer(files)` (:82) + `FixedLengthRecordReader` (:55), written down from the
TF-1.x semantics (the library is not installable here, see ref_numpy.py):
  * the filename queue yields the file list in a fresh random order every epoch;
  * ONE reader thread walks each file record by record and enqueues into a RandomShuffleQueue;
  * dequeue_many(batch) removes `batch` elements, each drawn uniformly from the queue's contents;
  * the reader refills the queue (this model: it is always ahead, so the queue is full at every dequeue).
The product path (data.BatchSource + csrc/shuffle_batch.cu) implements the same law with the same counter-based
random numbers; tests compare batches bit for bit.  Never imported by the product path.
"""
from __future__ import annotations

import numpy as np

MASK = (1 << 64) - 1


def splitmix64(seed, step, i):
    z = (seed + 0x9E3779B97F4A7C15 * (step * 65536 + i + 1)) & MASK
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK
    return z ^ (z >> 31)


class ShuffleQueue:
    """Pure-Python (exact integer arithmetic) reference of the dequeue / refill law."""

    def __init__(self, file_order, per_file, capacity=5384, batch=128, seed=0):
        self.file_order = [int(f) for f in file_order]
        self.per_file, self.capacity, self.batch, self.seed = per_file, capacity, batch, seed
        self.step = 0
        self.pos = capacity                     # the reader has already enqueued the first `capacity` records
        self.slots = [self.record_at(p) for p in range(capacity)]

    def record_at(self, p):
        return self.file_order[(p // self.per_file) % len(self.file_order)] * self.per_file + p % self.per_file

    def next_ids(self):
        s = self.slots
        for i in range(self.batch):
            u = splitmix64(self.seed, self.step, i) >> 32
            j = i + ((u * (self.capacity - i)) >> 32)
            s[i], s[j] = s[j], s[i]
        out = s[:self.batch]
        for i in range(self.batch):
            s[i] = self.record_at(self.pos + i)
        self.pos += self.batch
        self.step += 1
        return np.asarray(out, dtype=np.int64)
