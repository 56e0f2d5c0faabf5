"""Strided views (ref hw4code/strided_range_iterator.h:17-63: every `stride`-th element of a
range starting at `first`).  On tensors and arrays this is a slice; the helper exists so cipher
code reads like the reference's `strided_range(text + i, end, keyLength)`."""
from __future__ import annotations


def strided_range(seq, first: int, stride: int, last: int | None = None):
    if stride <= 0:
        raise ValueError("stride must be positive")
    return seq[first:last:stride]
