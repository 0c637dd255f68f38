"""Throughput / step-time counters and a JSON-lines writer (the reference only prints one line per epoch)."""
from __future__ import annotations

import json
import time
from typing import Optional

import torch


class StepMeter:
    """Device-timed throughput over windows of steps: CUDA events on the current stream (host clock on CPU)."""

    def __init__(self, device: torch.device, window: int = 10) -> None:
        self.cuda = device.type == "cuda"
        self.device, self.window = device, window
        self._count = 0
        self._items = 0
        self._start = None
        self.last = {}

    def _now(self):
        if self.cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record(torch.cuda.current_stream(self.device))
            return e
        return time.perf_counter()

    def step(self, items: int) -> Optional[dict]:
        if self._start is None:
            self._start = self._now()
            return None
        self._count += 1
        self._items += items
        if self._count < self.window:
            return None
        end = self._now()
        if self.cuda:
            end.synchronize()
            ms = self._start.elapsed_time(end)
        else:
            ms = (end - self._start) * 1e3
        self.last = {"steps": self._count, "ms_per_step": ms / self._count,
                     "items_per_sec": self._items / (ms / 1e3) if ms > 0 else float("inf")}
        self._count, self._items, self._start = 0, 0, end
        return self.last


class JsonlWriter:
    def __init__(self, path: Optional[str]) -> None:
        self.f = open(path, "a") if path else None

    def write(self, **record) -> None:
        if self.f:
            self.f.write(json.dumps(record) + "\n")
            self.f.flush()

    def close(self) -> None:
        if self.f:
            self.f.close()
            self.f = None
