"""Vigenere cipher workload (ref hw4code/create_cipher.cu:147-243, hw4code/solve_cipher.cu:42-182).

create: sanitise (lower-case, letters only) -> letter frequencies (GPU vs CPU check) -> random key
of `period` shifts in 1..25 -> encrypt with byte arithmetic (no mod-26 wrap).
solve: key length from the index of coincidence (first lag >= 4 with ioc > 1.6, confirmed at twice
that lag) -> one 256-bin histogram per key position (the text is never reordered) ->
most common byte is taken to be 'e' -> decrypt.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from ..ops import device as dev
from ..ops import host


def letter_frequency_cpu(raw: bytes | np.ndarray, top: int = 5):
    """CPU oracle: counts of lower-cased letters / letter total, top-`top` descending
    (ref hw4code/create_cipher.cu:69-96).  Returns (letters, frequencies)."""
    a = np.frombuffer(raw, dtype=np.uint8) if isinstance(raw, (bytes, bytearray)) else np.asarray(raw, np.uint8)
    low = np.where((a >= 65) & (a <= 90), a + 32, a)
    h = host.letter_histogram(low)[97:123].astype(np.float64)
    total = h.sum()
    order = np.argsort(-h, kind="stable")[:top]
    return [chr(97 + int(k)) for k in order], (h[order] / total).tolist()


def letter_frequency_gpu(clean: torch.Tensor, top: int = 5):
    """Integer counts on the device, one divide on the host: exact, where the reference summed
    1/len per letter in fp64 (ref hw4code/create_cipher.cu:99-145)."""
    h = dev.byte_histogram(clean).cpu().numpy()[97:123].astype(np.float64)
    order = np.argsort(-h, kind="stable")[:top]
    return [chr(97 + int(k)) for k in order], (h[order] / h.sum()).tolist()


def generate_key(period: int, seed: int = 123) -> np.ndarray:
    """`period` shifts in 1..25 (ref hw4code/create_cipher.cu:204-209 used thrust's
    default_random_engine(123); the stream differs, the distribution is the same)."""
    if period < 4:
        raise ValueError("Period must be at least 4!")      # ref create_cipher.cu:162-165
    rng = np.random.default_rng(seed)
    return rng.integers(1, 26, size=period, dtype=np.int32)


def key_string(shifts: np.ndarray) -> str:
    return "".join(chr(ord("a") + int(s) % 26) for s in shifts)


@dataclass
class CipherResult:
    clean: torch.Tensor
    cipher: torch.Tensor
    shifts: np.ndarray
    freq_gpu: tuple
    freq_cpu: tuple


def create_cipher(raw: bytes, period: int, seed: int = 123, wrap26: bool = False) -> CipherResult:
    d = torch.frombuffer(bytearray(raw), dtype=torch.uint8).cuda()
    buf, count = dev.sanitize_compact(d)
    clean = buf[: int(count.item())]
    fg = letter_frequency_gpu(clean)
    fc = letter_frequency_cpu(raw)
    shifts = generate_key(period, seed)
    cipher = dev.vigenere_shift(clean, torch.from_numpy(shifts).cuda(), wrap26=wrap26)
    return CipherResult(clean, cipher, shifts, fg, fc)


@dataclass
class SolveResult:
    key_length: int
    shifts: np.ndarray          # encryption shifts recovered (1..25)
    plain: torch.Tensor
    ioc: list


def find_key_length(cipher: torch.Tensor, max_lag: int = 512, chunk: int = 64, log=None):
    """ref hw4code/solve_cipher.cu:74-109, evaluating `chunk` lags per launch."""
    n = cipher.numel()
    key_length, lag, iocs = 0, 4, []
    counts, base = None, None
    while lag < min(max_lag, n):
        if counts is None or not (base <= lag < base + chunk):
            base = lag
            counts = dev.lag_match_counts(cipher, base, chunk).cpu().numpy()
        ioc = counts[lag - base] / ((n - lag) / 26.0)
        iocs.append((lag, float(ioc)))
        if log:
            log(f"Period {lag} ioc: {ioc}")
        if ioc > 1.6:
            if key_length == 0:
                key_length = lag
                lag = 2 * lag - 1
            elif 2 * key_length == lag:
                return key_length, iocs
            else:
                raise RuntimeError("Unusual pattern in text!")
        lag += 1
    raise RuntimeError("no key length found below max_lag")


def solve_cipher(cipher: torch.Tensor, wrap26: bool = False, log=None) -> SolveResult:
    key_length, iocs = find_key_length(cipher, log=log)
    hist = dev.strided_histograms(cipher, key_length).cpu().numpy()
    shifts = np.zeros(key_length, dtype=np.int32)
    for k in range(key_length):
        common = int(np.argmax(hist[k]))                       # assumed to be 'e'
        cipher_char = (common - ord("a")) % 256
        shifts[k] = (26 - 4 + cipher_char) % 26                # ref solve_cipher.cu:151-153
    plain = dev.vigenere_shift(cipher, torch.from_numpy(-shifts).cuda(), wrap26=wrap26)
    return SolveResult(key_length, shifts, plain, iocs)
