"""Tracing and metrics (the reference has neither: SURVEY §5.1, §5.5).

* ``nvtx_range(name)`` — NVTX range around a region when ``PSB_NVTX=1`` (Nsight Systems /
  Compute pick them up); a no-op otherwise, so the hot loop pays nothing by default.
* ``Metrics`` — counters and gauges kept by the engine (steps, pushes, pulls, bytes moved over
  NVLink by construction, last gradient norm) with a Prometheus text rendering and a JSON line
  logger for the per-step telemetry the reference only exposes as ``workers_received`` /
  ``total_workers`` on the wire.
"""
from __future__ import annotations

import contextlib
import json
import os
import sys
import time
from typing import Dict

import torch

_NVTX = os.environ.get("PSB_NVTX", "0") == "1"


@contextlib.contextmanager
def nvtx_range(name: str):
    if _NVTX and torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


class Metrics:
    def __init__(self, prefix: str = "psb"):
        self.prefix = prefix
        self.counters: Dict[str, float] = {}
        self.gauges: Dict[str, float] = {}
        self._t0 = time.time()

    def inc(self, name: str, value: float = 1.0) -> None:
        self.counters[name] = self.counters.get(name, 0.0) + value

    def set(self, name: str, value: float) -> None:
        self.gauges[name] = float(value)

    def snapshot(self) -> Dict[str, float]:
        out = dict(self.counters)
        out.update(self.gauges)
        out["uptime_s"] = time.time() - self._t0
        return out

    def prometheus(self) -> str:
        lines = []
        for k, v in sorted(self.counters.items()):
            lines += [f"# TYPE {self.prefix}_{k} counter", f"{self.prefix}_{k} {v}"]
        for k, v in sorted(self.gauges.items()):
            lines += [f"# TYPE {self.prefix}_{k} gauge", f"{self.prefix}_{k} {v}"]
        return "\n".join(lines) + "\n"

    def log_json(self, stream=sys.stderr, **extra) -> None:
        rec = self.snapshot()
        rec.update(extra)
        stream.write(json.dumps(rec) + "\n")
