"""Sample SM clocks and throttle reasons with nvidia-smi while a region is being timed
(B200_PROFILING.md "clocks DURING the timed region")."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
from typing import Any, Dict, List, Optional

_FIELDS = ["timestamp", "clocks.sm", "clocks.max.sm", "power.draw", "clocks_event_reasons.hw_slowdown",
           "clocks_event_reasons.hw_thermal_slowdown", "clocks_event_reasons.sw_thermal_slowdown",
           "clocks_event_reasons.sw_power_cap"]


class ClockSampler:
    """``with ClockSampler(gpu_index) as s: ...`` then ``s.summary()``."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 50) -> None:
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.window = None
        self.proc: Optional[subprocess.Popen] = None
        self.lines: List[str] = []
        self._thread: Optional[threading.Thread] = None

    def __enter__(self) -> "ClockSampler":
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        self.proc = subprocess.Popen(
            [exe, f"--query-gpu={','.join(_FIELDS)}", "--format=csv,noheader,nounits", "-lms", str(self.period_ms),
             "-i", str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)

        def pump() -> None:
            assert self.proc is not None and self.proc.stdout is not None
            for line in self.proc.stdout:
                self.lines.append(line.strip())

        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        # nvidia-smi takes a while to produce its first line (longer with 8 ranks starting one each);
        # wait for it so that the timed region is actually covered
        import time

        deadline = time.time() + 8.0
        while not self.lines and time.time() < deadline:
            time.sleep(0.02)
        return self

    def __exit__(self, *exc: Any) -> None:
        if self.proc is not None:
            self.proc.terminate()          # the exact child we started
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:  # pragma: no cover
                self.proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)

    def mark_start(self) -> None:
        """Start of the timed region (host clock); samples before it are dropped."""
        import datetime

        self.window = [datetime.datetime.now(), None]

    def mark_stop(self) -> None:
        import datetime

        if self.window is not None:
            self.window[1] = datetime.datetime.now()

    def summary(self) -> Dict[str, Any]:
        import datetime

        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) != len(_FIELDS):
                continue
            if self.window is not None:
                try:
                    ts = datetime.datetime.strptime(parts[0], "%Y/%m/%d %H:%M:%S.%f")
                except ValueError:
                    continue
                lo, hi = self.window
                if ts < lo or (hi is not None and ts > hi + datetime.timedelta(milliseconds=self.period_ms)):
                    continue
            parts = parts[1:]
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
                power.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power) if power else None}
