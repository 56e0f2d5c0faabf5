"""nvidia-smi clock / throttle sampler for benches (profiling recipe: sample DURING the timed
region, report the median SM clock and any slowdown reasons)."""
from __future__ import annotations

import shutil
import subprocess
import threading

_QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")
_NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu = gpu_index
        self.period = period_ms
        self.proc = None
        self.lines: list[str] = []
        self._thr = None
        self._mark = 0

    def start(self):
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        try:
            self.proc = subprocess.Popen(
                [exe, f"--id={self.gpu}", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits",
                 "-lms", str(self.period)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return self

        def pump():
            for line in self.proc.stdout:
                self.lines.append(line.strip())

        self._thr = threading.Thread(target=pump, daemon=True)
        self._thr.start()
        return self

    def mark(self) -> None:
        """Samples taken before this call (warm-up) are dropped unless nothing comes after."""
        self._mark = len(self.lines)

    def stop(self) -> dict:
        if self.proc is not None and not self.lines:
            import time

            for _ in range(100):           # a very short timed region: wait for the first sample
                if self.lines:
                    break
                time.sleep(0.02)
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            if self._thr is not None:
                self._thr.join(timeout=1)
        sm, mx, reasons = [], [], set()
        lines = self.lines[self._mark:] if len(self.lines) > self._mark else self.lines[-3:]
        for ln in lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, val in zip(_NAMES, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}
