"""Sample SM clocks / throttle reasons with nvidia-smi while a timed region runs
(B200_PROFILING.md, "clocks DURING the timed region")."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
import time

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    def __init__(self, gpu_index=0, period_ms=100):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.samples = []
        self._proc = None
        self._thread = None

    def start(self):
        if shutil.which("nvidia-smi") is None:
            return self
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + _QUERY,
                 "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self._proc = None
            return self

        def reader():
            for line in self._proc.stdout:
                parts = [p.strip() for p in line.split(",")]
                if len(parts) >= 8:
                    self.samples.append((time.time(), parts))

        self._thread = threading.Thread(target=reader, daemon=True)
        self._thread.start()
        return self

    def stop(self):
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=2)
            except Exception:
                self._proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        return self

    def summary(self, t0=None, t1=None):
        rows = [p for (t, p) in self.samples if (t0 is None or t >= t0 - 0.2) and (t1 is None or t <= t1 + 0.2)]
        if not rows:
            rows = [p for (_, p) in self.samples]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for p in rows:
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for nm, val in zip(names, p[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(rows)}
