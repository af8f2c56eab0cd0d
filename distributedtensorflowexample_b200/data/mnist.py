"""MNIST input pipeline with the TF1 tutorial loader's semantics (reference main.py:43-44,
worker.py:133,153; TF1 input_data.read_data_sets / DataSet.next_batch - SURVEY C2/T18/N13, Appendix D).

* idx.gz parsing (big-endian magic 0x803 images / 0x801 labels) - native reader in
  ``libdtfe_host.so`` with a pure-python fallback used only to cross-check it in tests.
* 60000 train -> first 5000 validation + 55000 train; 10000 test; images ``[N,784]`` float32 in
  [0,1]; labels one-hot float64 (``one_hot=True``) or uint8 ids.
* ``next_batch(B)``: shuffle at the first call and at every epoch boundary, per-process RNG;
  a batch that straddles an epoch end is ``rest of old epoch ++ head of new epoch``.
* No sharding across workers: every worker samples the full training set independently.

The reference checkout lacks ``train-images-idx3-ubyte.gz`` (its .MISSING_LARGE_BLOBS), and TF1
would download it; there is no network here.  Policy: use the train file if it is
present, otherwise split t10k 8000/2000 into train/test (labelled ``source='t10k-split'``),
otherwise fall back to synthetic data of the same shape.
"""
from __future__ import annotations

import gzip
import os
import struct
from dataclasses import dataclass
from pathlib import Path
from typing import Optional, Tuple

import numpy as np

IMAGE_MAGIC = 0x00000803
LABEL_MAGIC = 0x00000801
NUM_CLASSES = 10
IMAGE_PIXELS = 784

TRAIN_IMAGES = "train-images-idx3-ubyte.gz"
TRAIN_LABELS = "train-labels-idx1-ubyte.gz"
TEST_IMAGES = "t10k-images-idx3-ubyte.gz"
TEST_LABELS = "t10k-labels-idx1-ubyte.gz"


def read_idx_python(path: os.PathLike) -> np.ndarray:
    """Pure-python idx reader (oracle for the native one)."""
    with gzip.open(path, "rb") as f:
        raw = f.read()
    magic, = struct.unpack(">I", raw[:4])
    if magic == IMAGE_MAGIC:
        n, r, c = struct.unpack(">III", raw[4:16])
        data = np.frombuffer(raw, dtype=np.uint8, offset=16)
        if data.size != n * r * c:
            raise ValueError(f"{path}: payload {data.size} != {n}x{r}x{c}")
        return data.reshape(n, r, c)
    if magic == LABEL_MAGIC:
        n, = struct.unpack(">I", raw[4:8])
        data = np.frombuffer(raw, dtype=np.uint8, offset=8)
        if data.size != n:
            raise ValueError(f"{path}: payload {data.size} != {n}")
        return data
    raise ValueError(f"{path}: bad idx magic 0x{magic:08x}")


def read_idx(path: os.PathLike) -> np.ndarray:
    """idx.gz -> uint8 ndarray via the native reader; python fallback if the host lib is unavailable."""
    try:
        from ..utils import native_io
        return native_io.read_idx(str(path))
    except (ImportError, OSError, RuntimeError):
        return read_idx_python(path)


def dense_to_one_hot(labels: np.ndarray, num_classes: int = NUM_CLASSES) -> np.ndarray:
    out = np.zeros((labels.shape[0], num_classes), dtype=np.float64)   # float64 like TF1's numpy.zeros
    out[np.arange(labels.shape[0]), labels.astype(np.int64)] = 1.0
    return out


class DataSet:
    """TF1 ``DataSet``: in-memory arrays + epoch-shuffled ``next_batch``."""

    def __init__(self, images: np.ndarray, labels: np.ndarray, seed: Optional[int] = None, reshape: bool = True):
        if images.shape[0] != labels.shape[0]:
            raise ValueError(f"images {images.shape} vs labels {labels.shape}")
        if images.dtype == np.uint8:
            if reshape and images.ndim == 3:
                images = images.reshape(images.shape[0], -1)
            images = images.astype(np.float32) * np.float32(1.0 / 255.0)
        self._images = images
        self._labels = labels
        self._num_examples = images.shape[0]
        self._epochs_completed = 0
        self._index_in_epoch = 0
        self._rng = np.random.default_rng(seed)

    @property
    def images(self) -> np.ndarray:
        return self._images

    @property
    def labels(self) -> np.ndarray:
        return self._labels

    @property
    def num_examples(self) -> int:
        return self._num_examples

    @property
    def epochs_completed(self) -> int:
        return self._epochs_completed

    @property
    def label_ids(self) -> np.ndarray:
        """int32 class ids whatever the label encoding (the GPU path keeps ids, not one-hot)."""
        lab = self._labels
        return (lab.argmax(1) if lab.ndim == 2 else lab).astype(np.int32)

    def _shuffle(self) -> None:
        perm = self._rng.permutation(self._num_examples)
        self._images = self._images[perm]
        self._labels = self._labels[perm]

    def next_batch(self, batch_size: int, shuffle: bool = True) -> Tuple[np.ndarray, np.ndarray]:
        start = self._index_in_epoch
        if self._epochs_completed == 0 and start == 0 and shuffle:
            self._shuffle()
        if start + batch_size > self._num_examples:
            self._epochs_completed += 1
            rest = self._num_examples - start
            img_rest, lab_rest = self._images[start:], self._labels[start:]
            if shuffle:
                self._shuffle()
            new = batch_size - rest
            if new > self._num_examples:
                raise ValueError("batch_size larger than twice the dataset")
            self._index_in_epoch = new
            return (np.concatenate((img_rest, self._images[:new]), axis=0),
                    np.concatenate((lab_rest, self._labels[:new]), axis=0))
        self._index_in_epoch = start + batch_size
        return self._images[start:self._index_in_epoch], self._labels[start:self._index_in_epoch]


@dataclass
class Datasets:
    train: DataSet
    validation: DataSet
    test: DataSet
    source: str   # 'mnist' | 't10k-split' | 'synthetic'


def synthetic_arrays(n: int, seed: int = 0, num_classes: int = NUM_CLASSES, dim: int = IMAGE_PIXELS):
    """Learnable synthetic 784-dim data: each class has a fixed random prototype in [0,1] plus noise."""
    rng = np.random.default_rng(seed)
    protos = rng.random((num_classes, dim), dtype=np.float32)
    labels = rng.integers(0, num_classes, size=n).astype(np.uint8)
    x = protos[labels] * 0.6 + 0.4 * rng.random((n, dim), dtype=np.float32)
    return x.astype(np.float32), labels


def read_data_sets(train_dir: os.PathLike = "MNIST_data", one_hot: bool = False, validation_size: int = 5000,
                   seed: Optional[int] = None, allow_fallback: bool = True, synthetic: bool = False,
                   synthetic_train: int = 55000) -> Datasets:
    """``input_data.read_data_sets(train_dir, one_hot=...)`` (reference main.py:43-44)."""
    enc = (lambda l: dense_to_one_hot(l)) if one_hot else (lambda l: l)
    d = Path(train_dir)
    seeds = [None, None, None] if seed is None else [seed, seed + 1, seed + 2]
    if synthetic:
        xtr, ytr = synthetic_arrays(synthetic_train + validation_size, seed=0)
        xte, yte = synthetic_arrays(10000, seed=0)   # same prototypes (seed), different draw order is fine
        return Datasets(DataSet(xtr[validation_size:], enc(ytr[validation_size:]), seeds[0]),
                        DataSet(xtr[:validation_size], enc(ytr[:validation_size]), seeds[1]),
                        DataSet(xte, enc(yte), seeds[2]), "synthetic")
    have = {n: (d / n).exists() for n in (TRAIN_IMAGES, TRAIN_LABELS, TEST_IMAGES, TEST_LABELS)}
    if all(have.values()):
        tr_i, tr_l = read_idx(d / TRAIN_IMAGES), read_idx(d / TRAIN_LABELS)
        te_i, te_l = read_idx(d / TEST_IMAGES), read_idx(d / TEST_LABELS)
        if not 0 <= validation_size <= len(tr_i):
            raise ValueError(f"validation_size {validation_size} outside [0,{len(tr_i)}]")
        return Datasets(DataSet(tr_i[validation_size:], enc(tr_l[validation_size:]), seeds[0]),
                        DataSet(tr_i[:validation_size], enc(tr_l[:validation_size]), seeds[1]),
                        DataSet(te_i, enc(te_l), seeds[2]), "mnist")
    if not allow_fallback:
        missing = [n for n, ok in have.items() if not ok]
        raise FileNotFoundError(f"missing in {d}: {missing} (TF1 would download them; no network here)")
    if have[TEST_IMAGES] and have[TEST_LABELS]:
        te_i, te_l = read_idx(d / TEST_IMAGES), read_idx(d / TEST_LABELS)
        n_train = 8000
        n_val = min(validation_size, 500)
        return Datasets(DataSet(te_i[n_val:n_train], enc(te_l[n_val:n_train]), seeds[0]),
                        DataSet(te_i[:n_val], enc(te_l[:n_val]), seeds[1]),
                        DataSet(te_i[n_train:], enc(te_l[n_train:]), seeds[2]), "t10k-split")
    return read_data_sets(train_dir, one_hot, validation_size, seed, synthetic=True)


def find_data_dir(preferred: str = "MNIST_data") -> str:
    """Locate the idx files: the given dir, the repo's copy, or the mounted reference."""
    here = Path(__file__).resolve().parents[2]
    for cand in (Path(preferred), here / "MNIST_data", Path("/root/reference/MNIST_data")):
        if (cand / TEST_IMAGES).exists():
            return str(cand)
    return preferred
