"""Sample GPU clocks / throttle reasons with nvidia-smi while a timed region runs."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from typing import List, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")
_REASONS = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]


class ClockSampler:
    """``with ClockSampler(gpu_index) as c: ...; c.summary()`` -> {"sm_mhz", "sm_max_mhz", "reasons"}."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 200):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.proc: Optional[subprocess.Popen] = None
        self.lines: List[str] = []
        self.stamps: List[float] = []
        self.t0: Optional[float] = None
        self.t1: Optional[float] = None
        self.thread: Optional[threading.Thread] = None

    def mark_start(self) -> None:
        """Start of the timed region (the sampler itself should be started earlier: nvidia-smi needs a
        few hundred ms before its first line)."""
        self.t0 = time.time()

    def mark_end(self) -> None:
        self.t1 = time.time()

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i", str(self.gpu_index),
                 "-lms", str(self.period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return self

        def reader():
            for line in self.proc.stdout:
                self.lines.append(line.strip())
                self.stamps.append(time.time())

        self.thread = threading.Thread(target=reader, daemon=True)
        self.thread.start()
        return self

    def __exit__(self, *exc):
        if self.proc is not None:
            self.proc.terminate()          # the exact process we started
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        return False

    def summary(self) -> dict:
        sm, mx, power = [], [], []
        reasons = set()
        lines = self.lines
        if self.t0 is not None:
            t1 = self.t1 if self.t1 is not None else time.time()
            inside = [ln for ln, ts in zip(self.lines, self.stamps) if self.t0 <= ts <= t1 + 0.25]
            if inside:
                lines = inside
        for ln in lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
                power.append(float(parts[3]))
            except ValueError:
                continue
            for name, val in zip(_REASONS, parts[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power) if power else None}
