"""2-D heat-diffusion workload on one GPU (ref hw3code/gpuStencil.cu:87-140, 192-247, 281-308)."""
from __future__ import annotations

import numpy as np
import torch

from ..ops import device as dev
from ..utils.heat import SimParams


def heat_device(grid0: torch.Tensor, p: SimParams, variant: str = "shared", fused_bc: bool = True,
                iters: int | None = None) -> torch.Tensor:
    """Runs p.iters steps from grid0 (gy x gx float32 on the GPU) and returns the final grid.
    Per step: border decay next = curr*s (GPU semantics of the reference), stencil
    sweep, pointer swap.  With fused_bc the decay rides inside the sweep launch."""
    assert grid0.is_cuda and grid0.dtype == torch.float32 and grid0.shape == (p.gy, p.gx)
    cur = grid0.clone()
    nxt = grid0.clone()
    s = float(np.float32(p.bc_scale))
    n = p.iters if iters is None else iters
    for _ in range(n):
        if fused_bc:
            dev.heat_step(nxt, cur, p.order, p.xcfl, p.ycfl, variant, bc_scale=s)
        else:
            dev.heat_bc(nxt, cur, p.order, s)
            dev.heat_step(nxt, cur, p.order, p.xcfl, p.ycfl, variant)
        cur, nxt = nxt, cur
    return cur


class Grid:
    """Host + device copy of one (gy x gx) float32 grid (ref hw3code/Grid.{h,cu}): toGPU / fromGPU,
    pointer swap between two grids, and the text dump the reference calls saveStateToFile."""

    def __init__(self, gx: int, gy: int, host: np.ndarray | None = None):
        self.gx, self.gy = gx, gy
        self.h = np.zeros((gy, gx), dtype=np.float32) if host is None else np.ascontiguousarray(host, np.float32)
        self.d = torch.zeros((gy, gx), dtype=torch.float32, device="cuda")

    @classmethod
    def initial(cls, p: SimParams) -> "Grid":
        """sin(i dx) sin(j dy) initial condition (ref hw3code/main.cu:116-130), already on the GPU."""
        from ..ops import host as H

        g = cls(p.gx, p.gy, H.heat_init(p.gx, p.gy, p.dx, p.dy))
        g.to_gpu()
        return g

    def copy(self) -> "Grid":
        other = Grid(self.gx, self.gy, self.h.copy())
        other.d.copy_(self.d)
        return other

    def to_gpu(self) -> None:
        self.d.copy_(torch.from_numpy(self.h))

    def from_gpu(self) -> None:
        self.h = self.d.cpu().numpy()

    @staticmethod
    def swap(a: "Grid", b: "Grid") -> None:
        a.h, b.h = b.h, a.h
        a.d, b.d = b.d, a.d

    def save_state_to_file(self, identifier: str, directory: str = ".") -> str:
        """grid_<identifier>.txt, one grid row per line (ref hw3code/Grid.cu:77-101)."""
        from pathlib import Path

        from ..utils.checkpoint import save_grid_text

        path = str(Path(directory) / f"grid_{identifier}.txt")
        save_grid_text(torch.from_numpy(self.h), path)
        return path
