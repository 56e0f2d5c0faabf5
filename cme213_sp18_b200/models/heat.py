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
