"""Sample SM clocks / throttle reasons with ``nvidia-smi`` while a timed region runs
(the recipe in /opt/skills/guides/B200_PROFILING.md).  Used by bench.py and the benchmarks."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
from typing import Dict, List, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: int = 200) -> None:
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self.proc: Optional[subprocess.Popen] = None
        self.lines: List[str] = []
        self._thread: Optional[threading.Thread] = None

    def start(self) -> "ClockSampler":
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        self.proc = subprocess.Popen([exe, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits",
                                      "-i", str(self.gpu_index), "-lms", str(self.period_ms)],
                                     stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)

        def pump():
            assert self.proc is not None and self.proc.stdout is not None
            for line in self.proc.stdout:
                self.lines.append(line.strip())

        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> Dict:
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        return self.summary()

    def mark(self) -> int:
        """Position in the sample log; pass it to ``summary(since=...)`` to summarise only what came after."""
        return len(self.lines)

    def summary(self, since: int = 0) -> Dict:
        sm, mx, power = [], [], []
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines[since:]:
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2])); power.append(float(parts[3]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(power),
                "reasons": sorted(reasons), "samples": len(sm)}
