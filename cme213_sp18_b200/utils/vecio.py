"""Vector <-> text file helpers (ref hw1code/tests_q1.cpp:8-37, hw1code/tests_q2.cpp:35-78:
ReadVectorFromFile / WriteVectorToFile, one unsigned integer per line)."""
from __future__ import annotations

from pathlib import Path

import numpy as np


def read_vector_from_file(path) -> np.ndarray:
    vals = [int(tok) for tok in Path(path).read_text().split()]
    return np.array(vals, dtype=np.uint32)


def write_vector_to_file(path, vec) -> None:
    Path(path).parent.mkdir(parents=True, exist_ok=True)
    Path(path).write_text("".join(f"{int(v)}\n" for v in np.asarray(vec).ravel()))
