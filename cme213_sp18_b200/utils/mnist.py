"""MNIST-shaped data: IDX reader and a synthetic generator.

The reference reads the IDX files into doubles with raw 0..255 pixel values, one image per
column, no normalisation (ref fpcode/utils/mnist.cpp:9-66).  Here images stay uint8 [N, 784]
(the same bytes as the IDX payload) until the device converts them; labels are int32.
There is no network on the GPU box, so `synthetic_mnist` builds a learnable stand-in: every
class has a fixed random prototype image and samples are noisy copies of it.
"""
from __future__ import annotations

import struct
from pathlib import Path

import numpy as np
import torch

IMAGE_SIZE = 784
NUM_CLASSES = 10
NUM_TRAIN = 60000
NUM_TEST = 10000


def read_idx_images(path, max_images: int | None = None) -> torch.Tensor:
    """[N, rows*cols] uint8, read by the native host library. Header is big-endian (ref
    mnist.cpp reverse_int :9-18)."""
    from ..ops import host

    magic, (n, rows, cols) = host.idx_info(path)
    if magic != 2051:
        raise ValueError(f"{path}: bad IDX image magic {magic}")
    if max_images is not None:
        n = min(n, max_images)
    a = host.idx_read(path, 0, n, rows * cols)
    return torch.from_numpy(a.reshape(n, rows * cols))


def read_idx_labels(path, max_labels: int | None = None) -> torch.Tensor:
    from ..ops import host

    magic, (n, _, _) = host.idx_info(path)
    if magic != 2049:
        raise ValueError(f"{path}: bad IDX label magic {magic}")
    if max_labels is not None:
        n = min(n, max_labels)
    return torch.from_numpy(host.idx_read(path, 0, n, 1).astype(np.int32))


def write_idx_images(path, images_u8: torch.Tensor, rows: int = 28, cols: int = 28) -> None:
    a = images_u8.cpu().numpy().astype(np.uint8)
    Path(path).parent.mkdir(parents=True, exist_ok=True)
    with open(path, "wb") as f:
        f.write(struct.pack(">IIII", 2051, a.shape[0], rows, cols))
        f.write(a.tobytes())


def write_idx_labels(path, labels: torch.Tensor) -> None:
    a = labels.cpu().numpy().astype(np.uint8)
    Path(path).parent.mkdir(parents=True, exist_ok=True)
    with open(path, "wb") as f:
        f.write(struct.pack(">II", 2049, a.shape[0]))
        f.write(a.tobytes())


def synthetic_mnist(n: int, seed: int = 0, image_size: int = IMAGE_SIZE, num_classes: int = NUM_CLASSES,
                    pixel_max: int = 255, noise: float = 0.35):
    """(images uint8 [n, image_size] in 0..pixel_max, labels int32 [n]).  Raw 0..255 pixels
    saturate the sigmoids exactly as real MNIST does under the reference; pass
    pixel_max=1 for a well-conditioned problem."""
    g = torch.Generator().manual_seed(seed)
    proto = (torch.rand((num_classes, image_size), generator=g) < 0.25).float()
    labels = torch.randint(0, num_classes, (n,), generator=g, dtype=torch.int32)
    x = proto[labels.long()]
    flip = torch.rand((n, image_size), generator=g) < noise * 0.3
    x = torch.where(flip, 1.0 - x, x)
    x = x * (0.5 + 0.5 * torch.rand((n, image_size), generator=g))
    images = (x * pixel_max).round().clamp(0, 255).to(torch.uint8)
    return images, labels


def load_or_synthesize(data_dir, n_train: int = NUM_TRAIN, n_test: int = NUM_TEST, seed: int = 0,
                       pixel_max: int = 255):
    """The reference's three files if present under data_dir (ref fpcode/main.cpp:17-19),
    otherwise synthetic data of the same shapes.  Returns (x, labels, x_test, source)."""
    d = Path(data_dir) if data_dir else None
    if d is not None and (d / "train-images-idx3-ubyte").exists():
        x = read_idx_images(d / "train-images-idx3-ubyte", n_train)
        lab = read_idx_labels(d / "train-labels-idx1-ubyte", n_train)
        tpath = d / "t10k-images-idx3-ubyte"
        xt = read_idx_images(tpath, n_test) if tpath.exists() else synthetic_mnist(n_test, seed + 1)[0]
        return x, lab, xt, "idx"
    x, lab = synthetic_mnist(n_train, seed, pixel_max=pixel_max)
    xt, _ = synthetic_mnist(n_test, seed + 1, pixel_max=pixel_max)
    return x, lab, xt, "synthetic"
