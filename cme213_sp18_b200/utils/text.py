"""Text fixtures for the shift and cipher workloads.

The reference reads `mobydick.txt` (ref hw2code/main_q1.cu:91-118, hw4code/create_cipher.cu:168-177).
The GPU box has no datasets, so the workloads accept any file and fall back to synthetic
English-like text built here from a fixed paragraph with a seeded word shuffle.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np

_PARAGRAPH = (
    "The stencil sweeps the grid one row at a time while the halo rows travel to the neighbour "
    "over the switch fabric. Every thread keeps a small window of the column in registers and "
    "the centre row is shared through on chip memory, so each value is read from main memory "
    "once. When the sweep is done the edge rows are already in place on the other device and "
    "the next step can begin after a single flag is seen. The trainer works the same way: the "
    "gradient of each layer is summed across the devices by the kernel that also applies the "
    "weight decay and the learning step, and the fresh weights are written back to every "
    "replica before the next batch arrives. Nothing goes through the host, and nothing waits "
    "for a message that could have been a store. The letters of this text follow the usual "
    "English frequencies closely enough that the most common letter in every key position is "
    "the letter e, which is what the frequency attack expects to see there."
)


def synthetic_english(n_bytes: int, seed: int = 0) -> bytes:
    """Deterministic English-like text of exactly n_bytes (mixed case, punctuation, newlines)."""
    rng = np.random.default_rng(seed)
    sentences = [s.strip() + "." for s in _PARAGRAPH.split(".") if s.strip()]
    out, size = [], 0
    while size < n_bytes:
        order = rng.permutation(len(sentences))
        for k in order:
            s = sentences[k]
            sep = "\n" if rng.random() < 0.2 else " "
            out.append(s + sep)
            size += len(s) + 1
            if size >= n_bytes:
                break
    return "".join(out).encode("ascii")[:n_bytes]


def load_text(path: str | Path | None, fallback_bytes: int = 1_235_150, seed: int = 0) -> bytes:
    """File contents if the path exists, else synthetic text of the reference file's size."""
    if path is not None and Path(path).exists():
        return Path(path).read_bytes()
    return synthetic_english(fallback_bytes, seed)
