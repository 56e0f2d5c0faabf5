"""Weight export / import.

The reference has no resume; its only weight artefacts are Armadillo `raw_ascii` dumps of
W0/W1/b0/b1 (ref fpcode/neural_network.cpp:27-40, reloaded at :42-68).  This module writes and
reads that exact text layout (one matrix row per line, scientific notation) so dumps are
interchangeable, and adds a real checkpoint (`save_checkpoint` / `load_checkpoint`) carrying the
weights in the reference's W1/b1/W2/b2 layout plus the trainer position for resume.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch


def save_raw_ascii(mat: torch.Tensor, path) -> None:
    """Armadillo raw_ascii: rows on lines, values in %.16e-style scientific notation."""
    a = mat.detach().to(torch.float64).cpu().numpy()
    if a.ndim == 1:
        a = a.reshape(-1, 1)
    Path(path).parent.mkdir(parents=True, exist_ok=True)
    np.savetxt(path, a, fmt="%24.16e", delimiter=" ")


def load_raw_ascii(path) -> torch.Tensor:
    a = np.loadtxt(path, dtype=np.float64, ndmin=2)
    return torch.from_numpy(a)


def colmajor_bytes(mat: torch.Tensor) -> bytes:
    """The matrix as Armadillo holds it in memory (column-major fp64)."""
    return mat.detach().to(torch.float64).cpu().T.contiguous().numpy().tobytes()


def save_checkpoint(path, nn, epoch: int = 0, iteration: int = 0, extra: dict | None = None) -> None:
    Path(path).parent.mkdir(parents=True, exist_ok=True)
    torch.save({"format": "cme213b-mlp-v1", "model": nn.state_dict(), "epoch": epoch,
                "iteration": iteration, "extra": extra or {}}, path)


def load_checkpoint(path, nn) -> dict:
    ck = torch.load(path, map_location="cpu", weights_only=False)
    if ck.get("format") != "cme213b-mlp-v1":
        raise ValueError(f"{path}: not a cme213b checkpoint")
    nn.load_state_dict(ck["model"])
    return {"epoch": ck["epoch"], "iteration": ck["iteration"], "extra": ck["extra"]}


def save_grid_text(grid: torch.Tensor, path) -> None:
    """Text dump of a heat grid, one row per line (ref hw3code/Grid.cu:77-101 saveStateToFile)."""
    a = grid.detach().float().cpu().numpy()
    Path(path).parent.mkdir(parents=True, exist_ok=True)
    np.savetxt(path, a, fmt="%.8e")
