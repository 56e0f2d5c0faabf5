"""Timers (ref fpcode/inc/gpu_func.h:8-40 event_pair/start_timer/stop_timer; the reference
printed "seconds" for a millisecond value, SURVEY Q12 — units here are explicit)."""
from __future__ import annotations

import time
from contextlib import contextmanager

import torch


class EventTimer:
    """CUDA-event stopwatch on the current stream. elapsed_ms() synchronises on the stop event."""

    def __init__(self):
        self._start = torch.cuda.Event(enable_timing=True)
        self._stop = torch.cuda.Event(enable_timing=True)

    def start(self):
        self._start.record()
        return self

    def stop(self):
        self._stop.record()
        return self

    def elapsed_ms(self) -> float:
        self._stop.synchronize()
        return float(self._start.elapsed_time(self._stop))


def time_cuda_ms(fn, warmup: int = 3, iters: int = 10, flush_l2: bool = False) -> float:
    """Median device time of fn() in milliseconds (CUDA events, after warm-up)."""
    flush = None
    if flush_l2:
        flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    samples = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        t = EventTimer().start()
        fn()
        samples.append(t.stop().elapsed_ms())
    samples.sort()
    return samples[len(samples) // 2]


@contextmanager
def wall_timer(out: dict, key: str = "seconds"):
    t0 = time.perf_counter()
    try:
        yield
    finally:
        out[key] = time.perf_counter() - t0


def max_over_ranks_ms(ms: float) -> float:
    """Max of a per-rank device time over the process group (identity without one)."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return ms
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
