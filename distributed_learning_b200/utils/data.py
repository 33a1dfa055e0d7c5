"""Datasets, per-device partitioning, synthetic input and the host->device input pipeline.

Parity: reference ``src/data.py``.

* ``DataPartitioner`` / ``Partition`` (``src/data.py:10-41``): ``randperm(len)`` cut into
  ``total_dev`` equal parts of whole batches; partition index is
  ``node_id*node_dev + worker_id`` (``:64``).  As in the reference the permutation comes from
  the global torch RNG, which the training loop seeds (``src/main.py:31``).
* ``get_partition_mnist`` / ``get_partition_image_folder`` (``:84-127``): same folder layout
  (``R/mnist/MNIST/raw/*-ubyte[.gz]`` and ``R/ImageFolder/<class>/*.jpg``), same
  normalisation.  MNIST is read with a small IDX reader so nothing is downloaded.
* ``random_data_generator`` (``:129-132``): one real batch gives shape and label range, then
  random batches forever.  New: ``device=`` generates on the device; the
  reference generates on the host every step, which costs more than the training step.

New here: ``SyntheticDataset`` (shape-only stand-in when no dataset exists on disk and
``--random_input 1``), ``write_fake_mnist`` / ``write_fake_image_folder`` test fixtures, and
``PinnedPrefetcher``, the input pipeline for real end-to-end runs: batches are staged in
pinned host memory and copied on a side stream one step ahead of the compute stream.
"""
from __future__ import annotations

import functools
import gzip
import os
import struct
from typing import Callable, Dict, Iterator, Optional, Tuple

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset

PER_WORKER_BATCH_SIZE = 128

MNIST_MEAN, MNIST_STD = 0.1307, 0.3081
IMAGENET_MEAN, IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)

# model_type -> (C, H, W), classes
INPUT_SPECS: Dict[str, Tuple[Tuple[int, int, int], int]] = {
    "mnist": ((1, 28, 28), 10),
    "cifar": ((3, 32, 32), 10),
    "imagenet": ((3, 224, 224), 1000),
    "googlenet": ((3, 224, 224), 1000),
    "resnet50": ((3, 224, 224), 1000),
}


# ------------------------------------------------------------------------------------------
# partitioning
# ------------------------------------------------------------------------------------------
class Partition(Dataset):
    """View of ``data`` through a list of indices."""

    def __init__(self, data, index):
        self.data = data
        self.index = index

    def __len__(self):
        return len(self.index)

    def __getitem__(self, i):
        return self.data[int(self.index[i])]


class DataPartitioner:
    """Split ``data`` into ``total_dev`` disjoint partitions of a whole number of batches."""

    def __init__(self, data, total_dev: int, batch_size: int, generator: Optional[torch.Generator] = None):
        self.data = data
        data_len = len(data)
        batches_per_dev = data_len // total_dev // batch_size
        if batches_per_dev < 1:
            raise ValueError(
                f"dataset of {data_len} samples cannot give {total_dev} devices one batch of {batch_size} each")
        part_len = batches_per_dev * batch_size
        indexes = torch.randperm(data_len, generator=generator)
        self.partitions = [indexes[i * part_len:(i + 1) * part_len] for i in range(total_dev)]

    def use(self, partition: int) -> Partition:
        return Partition(self.data, self.partitions[partition])


def _get_partition_helper(node_id, worker_id, node_dev, total_dev, dataset, batch_size=PER_WORKER_BATCH_SIZE,
                          num_workers: int = 2) -> DataLoader:
    partition = DataPartitioner(dataset, total_dev, batch_size).use(node_id * node_dev + worker_id)
    return DataLoader(partition, batch_size=int(batch_size), shuffle=True, num_workers=num_workers)


# ------------------------------------------------------------------------------------------
# datasets
# ------------------------------------------------------------------------------------------
def _read_idx(path: str) -> np.ndarray:
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rb") as f:
        raw = f.read()
    zero, dtype_code, ndim = struct.unpack(">HBB", raw[:4])
    if zero != 0 or dtype_code != 0x08:
        raise ValueError(f"{path}: not an unsigned-byte IDX file")
    dims = struct.unpack(">" + "I" * ndim, raw[4:4 + 4 * ndim])
    return np.frombuffer(raw, dtype=np.uint8, offset=4 + 4 * ndim).reshape(dims)


class MNISTDataset(Dataset):
    """MNIST training split from IDX files under ``root/MNIST/raw`` (torchvision's layout)."""

    FILES = ("train-images-idx3-ubyte", "train-labels-idx1-ubyte")

    def __init__(self, root: str):
        raw = os.path.join(root, "MNIST", "raw")
        paths = []
        for name in self.FILES:
            for cand in (os.path.join(raw, name), os.path.join(raw, name + ".gz")):
                if os.path.exists(cand):
                    paths.append(cand)
                    break
            else:
                raise FileNotFoundError(f"{os.path.join(raw, name)}[.gz] not found (no download: this box is offline)")
        self.images = torch.from_numpy(_read_idx(paths[0]).copy())
        self.labels = torch.from_numpy(_read_idx(paths[1]).copy()).long()

    def __len__(self):
        return self.images.shape[0]

    def __getitem__(self, i):
        img = self.images[i].to(torch.float32).div_(255.0).sub_(MNIST_MEAN).div_(MNIST_STD).unsqueeze(0)
        return img, int(self.labels[i])


class SyntheticDataset(Dataset):
    """Random images of a model's input shape; used when no dataset is on disk."""

    def __init__(self, shape, num_classes: int, length: int = 8192, seed: int = 0):
        self.shape, self.num_classes, self.length, self.seed = tuple(shape), num_classes, length, seed

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        g = torch.Generator().manual_seed(self.seed * 1000003 + i)
        return torch.rand(self.shape, generator=g), int(torch.randint(self.num_classes, (1,), generator=g))


def _image_folder(root: str):
    import torchvision.datasets as datasets
    import torchvision.transforms as transforms

    return datasets.ImageFolder(
        os.path.join(root, "ImageFolder"),
        transform=transforms.Compose([
            transforms.Resize(256), transforms.CenterCrop(224), transforms.ToTensor(),
            transforms.Normalize(mean=IMAGENET_MEAN, std=IMAGENET_STD)]))


def _dataset_for(model_type: str, dataset_root: str, total_dev: int, batch_size: int):
    shape, classes = INPUT_SPECS[model_type]
    try:
        if model_type == "mnist":
            return MNISTDataset(os.path.join(dataset_root, "mnist"))
        if model_type in ("imagenet", "googlenet", "resnet50"):
            return _image_folder(dataset_root)
    except (FileNotFoundError, ImportError) as e:
        from .logging import Level, print_d

        print_d(f"dataset for '{model_type}' not found under {dataset_root} ({e}); using a synthetic stand-in", Level.WARNING)
    return SyntheticDataset(shape, classes, length=max(8192, total_dev * batch_size * 4))


def get_partition(model_type, node_id, worker_id, node_dev, total_dev, dataset_root,
                  batch_size=PER_WORKER_BATCH_SIZE, num_workers: int = 2) -> DataLoader:
    """This device's partition of ``model_type``'s dataset as a shuffling ``DataLoader``."""
    dataset = _dataset_for(model_type, dataset_root, total_dev, batch_size)
    return _get_partition_helper(node_id, worker_id, node_dev, total_dev, dataset, batch_size, num_workers)


# functools.partial of a module-level function: picklable, so a config can cross `spawn`
get_partition_mnist = functools.partial(get_partition, "mnist")
get_partition_cifar = functools.partial(get_partition, "cifar")
get_partition_image_folder = functools.partial(get_partition, "imagenet")

get_partition_dataset: Dict[str, Callable] = {
    "mnist": get_partition_mnist,
    "cifar": get_partition_cifar,
    "imagenet": get_partition_image_folder,
    "googlenet": get_partition_image_folder,
    "resnet50": get_partition_image_folder,
}


# ------------------------------------------------------------------------------------------
# synthetic input
# ------------------------------------------------------------------------------------------
def random_data_generator(data_pair, device=None) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
    """Endless random batches shaped like ``data_pair``.

    ``device=None`` reproduces the reference (host RNG every step).  With a device, the
    template is moved once and every batch is generated there, so ``get_data`` costs two
    kernel launches and ``data2dev`` nothing.
    """
    data, target = data_pair
    high = int(target.max()) + 1
    if device is not None:
        data, target = data.to(device), target.to(device)
    while True:
        yield torch.rand_like(data), torch.randint_like(target, high=high)


def synthetic_batches(model_type: str, batch_size: int, device, dtype=torch.float32):
    """Device-side random batches of a model's input shape, no dataset needed."""
    shape, classes = INPUT_SPECS[model_type]
    while True:
        yield (torch.rand((batch_size,) + shape, device=device, dtype=dtype),
               torch.randint(classes, (batch_size,), device=device))


class NativeHostRandom:
    """Fresh random batches made on the host by the native generator (``csrc/hostgen.cpp``).

    The job is the reference's ``random_data_generator`` on the host (``src/data.py:129-132``):
    every step a new ``(data, target)`` pair of the template's shape.  The reference draws it from
    torch's serial host RNG; here a few threads fill the *pinned staging slot itself*
    (``fill_into``, used by ``PinnedPrefetcher``), so a batch costs a few milliseconds and one pass
    over memory.  Iterating the object directly yields ordinary CPU tensors instead.
    ``dtype=torch.bfloat16`` writes bf16 on the host and halves the bytes of the H2D copy.
    """

    def __init__(self, shape, num_classes: int, seed: int = 0, dtype: torch.dtype = torch.float32,
                 nthreads: Optional[int] = None):
        self.shape = tuple(shape)               # (batch, C, H, W)
        self.num_classes = int(num_classes)
        self.seed = int(seed)
        self.dtype = dtype
        if nthreads is None:
            world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", 1)))
            nthreads = max(1, min(8, (os.cpu_count() or 1) // max(world, 1)))
        self.nthreads = int(os.environ.get("DLB200_HOSTGEN_THREADS", nthreads))
        self.batch_index = 0

    @classmethod
    def like(cls, data_pair, **kw) -> "NativeHostRandom":
        data, target = data_pair
        return cls(tuple(data.shape), int(target.max()) + 1, **kw)

    def template(self) -> Tuple[torch.Tensor, torch.Tensor]:
        return torch.empty(self.shape, dtype=self.dtype), torch.empty(self.shape[0], dtype=torch.int64)

    def fill_into(self, data: torch.Tensor, target: torch.Tensor) -> None:
        """Overwrite ``data`` / ``target`` (CPU tensors of the template's size) with the next batch.
        The memory format of ``data`` is irrelevant: the values are i.i.d."""
        from ..ops._lib import check, dtype_code, lib

        L = lib()
        check(L.dlb_host_fill_uniform(data.data_ptr(), dtype_code(data.dtype), data.numel(), self.seed, self.batch_index,
                                      self.nthreads))
        check(L.dlb_host_fill_randint(target.data_ptr(), target.numel(), self.num_classes, self.seed, self.batch_index))
        self.batch_index += 1

    def __iter__(self):
        return self

    def __next__(self):
        data, target = self.template()
        self.fill_into(data, target)
        return data, target


class PinnedPrefetcher:
    """Host->device input pipeline: a staging thread, pinned buffers, and a copy stream that runs
    one step ahead of the compute stream.

    ``source`` yields ``(data, target)`` CPU tensors.  A background thread pulls batches and
    stages them in one of ``depth`` pinned slots (batches that already live in pinned memory,
    e.g. from a ``DataLoader(pin_memory=True)``, are used where they are).  ``__next__`` makes
    the compute stream wait on the previous copy's event, hands out that batch and immediately
    issues the next host->device copy on ``copy_stream``, so the copy of step i+1 overlaps the
    compute of step i.  Bytes moved per step are in ``h2d_bytes_per_step``.  On a CPU device it
    degrades to a pass-through.
    """

    def __init__(self, source, device, depth: int = 3, data_dtype: Optional[torch.dtype] = None,
                 channels_last: bool = False):
        import queue
        import threading

        self.source = source if hasattr(source, "fill_into") else iter(source)
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.depth = max(2, depth)
        self.data_dtype = data_dtype
        self.channels_last = channels_last
        self.h2d_bytes_per_step = 0
        self._ready = None
        self._stop = False
        if self.cuda:
            self.copy_stream = torch.cuda.Stream(device=self.device)
            self._slots = [None] * self.depth
            self._free = queue.Queue()
            for i in range(self.depth):
                self._free.put(i)
            self._staged = queue.Queue(maxsize=self.depth)
            self._thread = threading.Thread(target=self._stage_loop, daemon=True, name="dlb200-prefetch")
            self._thread.start()
        self._queue_next()

    # -- staging thread ------------------------------------------------------------------------
    def _stage_loop(self):
        try:
            if hasattr(self.source, "fill_into"):
                self._fill_loop()
                return
            for data, target in self.source:
                if self._stop:
                    break
                if data.is_pinned() and target.is_pinned():
                    item = (data, target, None)
                else:
                    i = self._free.get()
                    if self._slots[i] is None:
                        fmt = torch.channels_last if (self.channels_last and data.dim() == 4) else torch.contiguous_format
                        self._slots[i] = (torch.empty(data.shape, dtype=data.dtype).contiguous(memory_format=fmt).pin_memory(),
                                          torch.empty_like(target).pin_memory(), torch.cuda.Event())
                    hd, ht, free_ev = self._slots[i]
                    free_ev.synchronize()          # the copy that last read this slot has finished
                    hd.copy_(data)
                    ht.copy_(target)
                    item = (hd, ht, i)
                self._staged.put(item)
            self._staged.put(None)
        except BaseException as e:  # noqa: BLE001 - re-raised in the consumer
            self._staged.put(e)

    def _fill_loop(self):
        """Sources that can write a batch in place (``NativeHostRandom``): no intermediate tensor,
        the generator's threads write the pinned slot the copy engine reads."""
        src = self.source
        while not self._stop:
            i = self._free.get()
            if self._stop:
                break
            if self._slots[i] is None:
                data, target = src.template()
                fmt = torch.channels_last if (self.channels_last and data.dim() == 4) else torch.contiguous_format
                self._slots[i] = (data.contiguous(memory_format=fmt).pin_memory(), target.pin_memory(), torch.cuda.Event())
            hd, ht, free_ev = self._slots[i]
            free_ev.synchronize()
            src.fill_into(hd, ht)
            self._staged.put((hd, ht, i))

    # -- consumer side ---------------------------------------------------------------------------
    def _queue_next(self):
        if not self.cuda:
            nx = next(self.source, None)
            self._ready = None if nx is None else (nx[0], nx[1], None)
            return
        item = self._staged.get()
        if item is None:
            self._ready = None
            return
        if isinstance(item, BaseException):
            raise item
        hd, ht, slot = item
        with torch.cuda.stream(self.copy_stream):
            dd = hd.to(self.device, non_blocking=True)
            dt = ht.to(self.device, non_blocking=True)
            if self.data_dtype is not None and dd.dtype != self.data_dtype:
                dd = dd.to(self.data_dtype)
            done = torch.cuda.Event()
            done.record(self.copy_stream)
            if slot is not None:
                self._slots[slot][2].record(self.copy_stream)
        if slot is not None:
            self._free.put(slot)
        self.h2d_bytes_per_step = hd.numel() * hd.element_size() + ht.numel() * ht.element_size()
        self._ready = (dd, dt, done, (hd, ht))     # keep the host tensors alive until the copy is consumed

    def __iter__(self):
        return self

    def __next__(self):
        if self._ready is None:
            raise StopIteration
        data, target, done = self._ready[:3]
        if done is not None:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(done)
            data.record_stream(cur)
            target.record_stream(cur)
        self._queue_next()
        return data, target

    def close(self) -> None:
        self._stop = True
        if self.cuda:
            self._free.put(0)          # wake a staging thread that waits for a free slot


# ------------------------------------------------------------------------------------------
# fixtures (no network on the build or GPU boxes)
# ------------------------------------------------------------------------------------------
def write_fake_mnist(root: str, n: int = 2048, seed: int = 0) -> str:
    """Write a random MNIST-format training split under ``root/mnist/MNIST/raw``; returns ``root``."""
    raw = os.path.join(root, "mnist", "MNIST", "raw")
    os.makedirs(raw, exist_ok=True)
    rng = np.random.RandomState(seed)
    images = rng.randint(0, 256, size=(n, 28, 28), dtype=np.uint8)
    labels = rng.randint(0, 10, size=(n,), dtype=np.uint8)
    with open(os.path.join(raw, "train-images-idx3-ubyte"), "wb") as f:
        f.write(struct.pack(">HBBIII", 0, 0x08, 3, n, 28, 28) + images.tobytes())
    with open(os.path.join(raw, "train-labels-idx1-ubyte"), "wb") as f:
        f.write(struct.pack(">HBBI", 0, 0x08, 1, n) + labels.tobytes())
    # torchvision's MNIST also wants the test split to consider the dataset present
    with open(os.path.join(raw, "t10k-images-idx3-ubyte"), "wb") as f:
        f.write(struct.pack(">HBBIII", 0, 0x08, 3, 16, 28, 28) + images[:16].tobytes())
    with open(os.path.join(raw, "t10k-labels-idx1-ubyte"), "wb") as f:
        f.write(struct.pack(">HBBI", 0, 0x08, 1, 16) + labels[:16].tobytes())
    return root


def write_fake_image_folder(root: str, n: int = 256, classes: int = 4, size: int = 64, seed: int = 0) -> str:
    """Write ``n`` random JPEGs in ``classes`` class folders under ``root/ImageFolder``."""
    from PIL import Image

    rng = np.random.RandomState(seed)
    for i in range(n):
        d = os.path.join(root, "ImageFolder", f"class{i % classes:03d}")
        os.makedirs(d, exist_ok=True)
        arr = rng.randint(0, 256, size=(size, size, 3), dtype=np.uint8)
        Image.fromarray(arr).save(os.path.join(d, f"img{i:05d}.jpg"), quality=60)
    return root


# ------------------------------------------------------------------------------------------
# legacy API: all of this node's partitions at once
# ------------------------------------------------------------------------------------------
def _partition_helper(node_dev: int, total_dev: int, dataset, batch_size: int = PER_WORKER_BATCH_SIZE, node_id=None):
    """Loaders for every device of this node (reference ``_partition_helper``, ``src/data.py:44-56``;
    unused by the reference's own experiments).  The node index comes from ``dist.get_rank()``
    unless ``node_id`` is given."""
    if node_id is None:
        import torch.distributed as dist

        node_id = dist.get_rank()
    partitioner = DataPartitioner(dataset, total_dev, batch_size)
    loaders = [DataLoader(partitioner.use(node_id * node_dev + i), batch_size=int(batch_size), shuffle=True)
               for i in range(node_dev)]
    return loaders, batch_size


def partition_mnist(node_dev: int, total_dev: int, dataset_root: str, batch_size: int = PER_WORKER_BATCH_SIZE, node_id=None):
    """Reference ``partition_mnist`` (``src/data.py:71-82``)."""
    return _partition_helper(node_dev, total_dev, _dataset_for("mnist", dataset_root, total_dev, batch_size), batch_size, node_id)


def partition_image_folder(node_dev: int, total_dev: int, dataset_root: str, batch_size: int = PER_WORKER_BATCH_SIZE, node_id=None):
    """Reference ``partition_image_folder`` (``src/data.py:97-111``)."""
    return _partition_helper(node_dev, total_dev, _dataset_for("imagenet", dataset_root, total_dev, batch_size), batch_size, node_id)
