"""Span tracing: NVTX ranges for a profiler and a Chrome-trace file for everything else.

The reference has stopwatches only (ref fpcode/inc/gpu_func.h:8-40, hw2code/common/mp1-util.h:1-35)
and no per-phase or communication timing.  A `Tracer` records named spans; on a GPU a
span is a pair of CUDA events on the current stream (device time, no synchronisation while
recording), on the CPU it is a pair of perf_counter readings.  `export_chrome_trace` writes the
spans in the Trace Event format (chrome://tracing, Perfetto), one row per rank.

    tr = Tracer(rank=info.rank)
    with tr.span("epoch", epoch=0):
        ...
    tr.export_chrome_trace("trace.json")

Set CME_NVTX=1 to also emit NVTX ranges (visible to nsys / ncu --nvtx).
"""
from __future__ import annotations

import json
import os
import time
from contextlib import contextmanager
from pathlib import Path

import torch

_NVTX = os.environ.get("CME_NVTX", "0") not in ("", "0")


@contextmanager
def nvtx_range(name: str):
    """NVTX push/pop when CME_NVTX is set and CUDA is up; otherwise free."""
    on = _NVTX and torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


class Tracer:
    def __init__(self, rank: int = 0, device: str | None = None, enabled: bool = True):
        self.rank = rank
        self.enabled = enabled
        self.cuda = (device or ("cuda" if torch.cuda.is_available() else "cpu")).startswith("cuda")
        self._t0 = time.perf_counter()
        self._origin = None            # CUDA event all device spans are measured against
        self._spans = []               # (name, start, stop, args); events on CUDA, seconds on the CPU
        if self.cuda and enabled:
            self._origin = torch.cuda.Event(enable_timing=True)
            self._origin.record()

    @contextmanager
    def span(self, name: str, **args):
        if not self.enabled:
            yield
            return
        with nvtx_range(name):
            if self.cuda:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                try:
                    yield
                finally:
                    b.record()
                    self._spans.append((name, a, b, args))
            else:
                a = time.perf_counter()
                try:
                    yield
                finally:
                    self._spans.append((name, a, time.perf_counter(), args))

    def spans_ms(self) -> list[tuple[str, float, float, dict]]:
        """(name, start_ms, duration_ms, args) with start relative to the tracer's creation."""
        out = []
        if self.cuda:
            torch.cuda.synchronize()
            for name, a, b, args in self._spans:
                out.append((name, self._origin.elapsed_time(a), a.elapsed_time(b), args))
        else:
            for name, a, b, args in self._spans:
                out.append((name, (a - self._t0) * 1e3, (b - a) * 1e3, args))
        return out

    def totals_ms(self) -> dict[str, float]:
        tot: dict[str, float] = {}
        for name, _, dur, _ in self.spans_ms():
            tot[name] = tot.get(name, 0.0) + dur
        return tot

    def export_chrome_trace(self, path) -> Path:
        events = [{"name": "process_name", "ph": "M", "pid": self.rank, "args": {"name": f"rank {self.rank}"}}]
        for name, start, dur, args in self.spans_ms():
            events.append({"name": name, "ph": "X", "pid": self.rank, "tid": 0, "ts": start * 1e3,
                           "dur": dur * 1e3, "args": args})
        path = Path(path)
        path.parent.mkdir(parents=True, exist_ok=True)
        path.write_text(json.dumps({"traceEvents": events, "displayTimeUnit": "ms"}))
        return path
