"""Data pipeline: MNIST-shaped uint8 dataset + per-rank sharded batches.

Covers reference data.py:6-27 (torchvision MNIST + ToTensor + DistributedSampler + DataLoader)
without torchvision, PIL or worker processes:

* the whole dataset is one uint8 tensor (47 MB for MNIST) read straight from the IDX files
  torchvision would have downloaded to ``<root>/MNIST/raw`` — or, offline, a deterministic
  synthetic stand-in of the same shape (BASELINE.json);
* on CUDA the dataset is resident on the device and a batch is a gather + u8->f32 scale kernel
  (ToTensor semantics: x/255, shape [b,1,28,28]; labels int64);
* batches follow DistributedSampler order exactly, last short batch kept (drop_last=False).
"""
from __future__ import annotations

import gzip
import os
import struct
from dataclasses import dataclass

import torch

from ..parallel.sampler import ShardedSampler

_IDX_FILES = {
    True: ("train-images-idx3-ubyte", "train-labels-idx1-ubyte"),
    False: ("t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte"),
}


@dataclass
class ImageDataset:
    images: torch.Tensor  # uint8 [N, H, W]
    labels: torch.Tensor  # int64 [N]
    source: str = "synthetic"

    def __len__(self) -> int:
        return int(self.images.shape[0])


# ----------------------------------------------------------------------------- IDX format
def read_idx(path: str) -> torch.Tensor:
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rb") as f:
        raw = f.read()
    zero, dtype_code, ndim = struct.unpack(">HBB", raw[:4])
    if zero != 0 or dtype_code != 0x08:
        raise ValueError(f"{path}: not a uint8 IDX file (magic {raw[:4].hex()})")
    dims = struct.unpack(">" + "I" * ndim, raw[4:4 + 4 * ndim])
    data = torch.frombuffer(bytearray(raw[4 + 4 * ndim:]), dtype=torch.uint8)
    return data.reshape(dims).clone()


def write_idx(path: str, t: torch.Tensor) -> None:
    t = t.to(torch.uint8).contiguous()
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "wb") as f:
        f.write(struct.pack(">HBB", 0, 0x08, t.dim()))
        f.write(struct.pack(">" + "I" * t.dim(), *t.shape))
        f.write(t.numpy().tobytes())


def write_mnist_idx(root: str, ds: ImageDataset, train: bool = True) -> None:
    """Write ``ds`` where torchvision's MNIST(root=...) looks for its raw files."""
    img_name, lab_name = _IDX_FILES[train]
    raw = os.path.join(root, "MNIST", "raw")
    write_idx(os.path.join(raw, img_name), ds.images)
    write_idx(os.path.join(raw, lab_name), ds.labels.to(torch.uint8))


# ----------------------------------------------------------------------------- datasets
def synthetic_mnist(n: int = 60000, seed: int = 0, h: int = 28, w: int = 28,
                    num_classes: int = 10) -> ImageDataset:
    """Deterministic MNIST-shaped data that is actually learnable: each class is a fixed smooth
    template, each sample is its class template under random shift, gain and pixel noise."""
    g = torch.Generator().manual_seed(seed)
    coarse = torch.rand(num_classes, 1, 7, 7, generator=g)
    templates = torch.nn.functional.interpolate(coarse, size=(h, w), mode="bilinear",
                                                align_corners=False)[:, 0]
    templates = (templates - templates.amin(dim=(1, 2), keepdim=True))
    templates = templates / templates.amax(dim=(1, 2), keepdim=True)
    labels = torch.randint(0, num_classes, (n,), generator=g)
    gain = 0.6 + 0.4 * torch.rand(n, 1, 1, generator=g)
    images = torch.empty(n, h, w, dtype=torch.uint8)
    chunk = 8192
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        t = templates[labels[s:e]] * gain[s:e]
        dx = int(torch.randint(-2, 3, (1,), generator=g))
        dy = int(torch.randint(-2, 3, (1,), generator=g))
        t = torch.roll(t, shifts=(dy, dx), dims=(1, 2))
        t = t + 0.15 * torch.randn(e - s, h, w, generator=g)
        images[s:e] = (t.clamp(0, 1) * 255).to(torch.uint8)
    return ImageDataset(images, labels.to(torch.int64), source=f"synthetic(seed={seed})")


def _find_idx(raw: str, train: bool):
    img_name, lab_name = _IDX_FILES[train]
    for ext in ("", ".gz"):
        ip, lp = os.path.join(raw, img_name + ext), os.path.join(raw, lab_name + ext)
        if os.path.exists(ip) and os.path.exists(lp):
            return ip, lp
    return None


def _try_download(root: str, train: bool) -> str | None:
    """What the reference does (reference data.py:11-14: ``datasets.MNIST(..., download=True)``):
    let torchvision fetch the files into ``<root>/MNIST/raw``.  Returns None on success, else the
    reason (no torchvision, no network, ...).  DDP_B200_NO_DOWNLOAD=1 skips the attempt."""
    if os.environ.get("DDP_B200_NO_DOWNLOAD", "0") == "1":
        return "download disabled by DDP_B200_NO_DOWNLOAD=1"
    try:
        import socket
        from torchvision import datasets
        prev = socket.getdefaulttimeout()
        socket.setdefaulttimeout(float(os.environ.get("DDP_B200_DOWNLOAD_TIMEOUT", "10")))
        try:
            datasets.MNIST(root=root, train=train, download=True)
        finally:
            socket.setdefaulttimeout(prev)
        return None
    except Exception as e:  # ImportError, URLError, RuntimeError("Error downloading ...")
        return f"{type(e).__name__}: {str(e).splitlines()[0][:160] if str(e) else ''}"


def load_mnist(root: str = "./data", train: bool = True, synthetic_fallback: bool = True,
               synthetic_size: int | None = None, download: bool = True, verbose: bool = True) -> ImageDataset:
    """MNIST for training, in the order the reference would get it (reference data.py:11-14):

    1. the IDX files torchvision keeps under ``<root>/MNIST/raw`` (plain or .gz), if present;
    2. else, with ``download`` (default), torchvision's own download into that directory;
    3. else — only with ``synthetic_fallback`` — a deterministic synthetic stand-in of the same
       shape, announced with a loud warning (it is NOT MNIST; it exists so that the offline
       benchmark boxes and the tests can run: BASELINE.json).

    ``verbose`` prints which source was used (``ImageDataset.source`` records it either way)."""
    import warnings
    raw = os.path.join(root, "MNIST", "raw")
    found = _find_idx(raw, train)
    why = None
    if found is None and download:
        why = _try_download(root, train)
        if why is None:
            found = _find_idx(raw, train)
            why = None if found else "torchvision reported success but the IDX files are missing"
    if found is not None:
        ip, lp = found
        ds = ImageDataset(read_idx(ip), read_idx(lp).to(torch.int64), source=ip)
        if verbose:
            print(f"[data] MNIST {'train' if train else 'test'} set: {len(ds)} images from {ip}", flush=True)
        return ds
    if not synthetic_fallback:
        raise FileNotFoundError(f"no MNIST IDX files under {raw} and they could not be downloaded ({why})")
    n = synthetic_size or int(os.environ.get("DDP_B200_SYNTHETIC_SIZE", 60000 if train else 10000))
    ds = synthetic_mnist(n=n, seed=0 if train else 1)
    msg = (f"no MNIST files under {raw} and no download ({why}): training on {n} SYNTHETIC MNIST-shaped "
           f"images ({ds.source}), not on MNIST.  Put the four IDX files there, or pass "
           f"synthetic_fallback=False to make this an error.")
    warnings.warn(msg, RuntimeWarning, stacklevel=2)
    if verbose:
        print(f"[data] WARNING: {msg}", flush=True)
    return ds


# ----------------------------------------------------------------------------- batch loaders
class BatchLoader:
    """Iterates this rank's batches for the sampler's current epoch: (images, labels) with
    images float32 [b,1,H,W] in [0,1] and labels int64 [b] on ``device``."""

    def __init__(self, dataset: ImageDataset, sampler: ShardedSampler, batch_size: int,
                 device: torch.device | str = "cpu", target=None):
        #: optional callable returning (x_buf, y_buf) device buffers to gather into (the engine's
        #: static graph inputs), saving a device-to-device copy per step
        self.target = target
        self.dataset = dataset
        self.sampler = sampler
        self.batch_size = int(batch_size)
        self.device = torch.device(device)
        self.on_cuda = self.device.type == "cuda"
        H, W = dataset.images.shape[1:]
        if self.on_cuda:
            from ..ops import native as nat
            self._nat = nat
            nat.lib()
            self._images = dataset.images.to(self.device)
            self._labels = dataset.labels.to(self.device)
            self._x = [torch.empty(self.batch_size, 1, H, W, device=self.device) for _ in range(2)]
            self._y = [torch.empty(self.batch_size, dtype=torch.int64, device=self.device)
                       for _ in range(2)]
        else:
            self._images = dataset.images
            self._labels = dataset.labels

    def __len__(self) -> int:
        return (len(self.sampler) + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        idx = self.sampler.indices()
        if self.on_cuda:
            idx = idx.to(self.device)
        for i, s in enumerate(range(0, idx.numel(), self.batch_size)):
            sel = idx[s:s + self.batch_size]
            b = sel.numel()
            if self.on_cuda:
                xb, yb = self.target() if self.target is not None else (self._x[i & 1], self._y[i & 1])
                x, y = xb[:b], yb[:b]
                self._nat.gather_u8(self._images, self._labels, sel, x, y)
                yield x, y
            else:
                x = self._images[sel].to(torch.float32).div_(255.0).unsqueeze(1)
                yield x, self._labels[sel]


def get_dataloader(batch_size: int, world_size: int, rank: int, device="cpu", root: str = "./data",
                   dataset: ImageDataset | None = None, target=None):
    """Same call shape and return value as the reference (reference data.py:6,27):
    ``(dataloader, sampler)``; call ``sampler.set_epoch(e)`` before iterating epoch ``e``."""
    # rank 0 announces the source; the other ranks load (or download) quietly
    ds = dataset if dataset is not None else load_mnist(root=root, train=True, verbose=(rank == 0))
    sampler = ShardedSampler(len(ds), num_replicas=world_size, rank=rank, shuffle=True)
    return BatchLoader(ds, sampler, batch_size, device=device, target=target), sampler
