"""ULP comparison helpers (ref fpcode/utils/test_utils.h:6-81 almost_equal_ulps /
almost_equal_matrix, hw2code/common/mp1-util.h:47-69 AlmostEqualUlps).  The distance itself is
computed by the native host library (`ops.host.max_ulp`)."""
from __future__ import annotations

import numpy as np

from ..ops import host


def almost_equal_ulps(a: float, b: float, max_ulps: int = 512, dtype=np.float64) -> bool:
    """Scalars within max_ulps units in the last place (512 fp64 in the reference's fpcode)."""
    d, _ = host.max_ulp(np.array([a], dtype=dtype), np.array([b], dtype=dtype))
    return d <= max_ulps


def almost_equal_matrix(m, raw, max_ulps: int = 512, column_major: bool = True) -> bool:
    """A 2-D array against a flat buffer in column- or row-major order (the reference compares an
    arma::mat with a raw device buffer this way)."""
    m = np.asarray(m)
    flat = np.asarray(raw, dtype=m.dtype).reshape(-1)
    other = flat.reshape(m.shape[::-1]).T if column_major else flat.reshape(m.shape)
    d, _ = host.max_ulp(np.ascontiguousarray(m), np.ascontiguousarray(other))
    return d <= max_ulps


def max_ulps_f32(a, b) -> int:
    """Largest fp32 ULP distance of two arrays (100 for PageRank, 512 for the stencil)."""
    return host.max_ulp(np.asarray(a, np.float32), np.asarray(b, np.float32))[0]
