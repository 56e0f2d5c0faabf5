"""Simulation parameters of the 2-D heat-diffusion workload (ref hw3code/simParams.{h,cpp})."""
from __future__ import annotations

import math
from dataclasses import dataclass
from pathlib import Path


@dataclass
class SimParams:
    """Same inputs as the reference's params.in (ref hw3code/simParams.cpp:7-51): interior size,
    physical lengths, iteration count and stencil order; everything else is derived."""

    nx: int = 4096
    ny: int = 4096
    lx: float = 1.0
    ly: float = 1.0
    iters: int = 400
    order: int = 8

    def __post_init__(self):
        if self.order not in (2, 4, 8):
            raise ValueError("order must be 2, 4 or 8")
        if self.nx <= 0 or self.ny <= 0 or self.lx <= 0 or self.ly <= 0 or self.iters < 0:
            raise ValueError("nx, ny, lx, ly must be positive and iters non-negative")

    @classmethod
    def from_file(cls, path: str | Path) -> "SimParams":
        """Whitespace-separated `nx ny / lx ly / iters / order` (ref hw3code/params.in)."""
        tok = Path(path).read_text().split()
        if len(tok) < 6:
            raise ValueError(f"{path}: expected 6 numbers (nx ny lx ly iters order)")
        return cls(int(tok[0]), int(tok[1]), float(tok[2]), float(tok[3]), int(tok[4]), int(tok[5]))

    def to_file(self, path: str | Path) -> None:
        Path(path).write_text(f"{self.nx} {self.ny}\n{self.lx:g} {self.ly:g}\n{self.iters}\n{self.order}\n")

    @property
    def border(self) -> int:
        return self.order // 2

    @property
    def gx(self) -> int:
        return self.nx + 2 * self.border

    @property
    def gy(self) -> int:
        return self.ny + 2 * self.border

    @property
    def dx(self) -> float:
        return self.lx / (self.gx - 1)

    @property
    def dy(self) -> float:
        return self.ly / (self.gy - 1)

    @property
    def dt(self) -> float:
        # ref hw3code/simParams.cpp:53-77: just under the stability limit for each order
        dx2, dy2 = self.dx * self.dx, self.dy * self.dy
        if self.order == 2:
            return (0.5 - 0.01) * (dx2 * dy2) / (dx2 + dy2)
        if self.order == 4:
            return (0.5 - 0.01) * (12 * dx2 * dy2) / (16 * (dx2 + dy2))
        return (0.5 - 0.01) * (5040 * dx2 * dy2) / (8064 * (dx2 + dy2))

    @property
    def _scale(self) -> float:
        return {2: 1.0, 4: 12.0, 8: 5040.0}[self.order]

    @property
    def xcfl(self) -> float:
        return self.dt / (self._scale * self.dx * self.dx)

    @property
    def ycfl(self) -> float:
        return self.dt / (self._scale * self.dy * self.dy)

    @property
    def bc_scale(self) -> float:
        """Border decay per step, exp(-2 dt) (ref hw3code/BC.h:62-63)."""
        return math.exp(-2.0 * self.dt)

    def calc_bytes(self) -> int:
        """The reference's byte model: iters*nx*ny*{6,10,18}*4 (ref hw3code/simParams.cpp:79-93)."""
        words = {2: 6, 4: 10, 8: 18}[self.order]
        return self.iters * self.nx * self.ny * words * 4

    def real_bytes(self) -> int:
        """Bytes that must cross HBM: one read and one write of the interior per step."""
        return self.iters * self.nx * self.ny * 8
