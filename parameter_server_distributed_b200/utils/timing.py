"""Device timing hygiene (B200_PROFILING.md): CUDA events, L2 flush, clock sampling."""
from __future__ import annotations

import statistics
import subprocess
import threading
from typing import Callable, List, Optional

import torch

_L2_FLUSH = {}


def flush_l2(device) -> None:
    """Overwrite a buffer larger than the 126 MB L2 so the next kernel starts cold."""
    device = torch.device(device)
    buf = _L2_FLUSH.get(device)
    if buf is None:
        buf = torch.empty(256 << 20, dtype=torch.uint8, device=device)
        _L2_FLUSH[device] = buf
    buf.zero_()


def cuda_time_ms(fn: Callable[[], None], iters: int = 10, warmup: int = 3,
                 flush: bool = True, device=None) -> List[float]:
    """Per-iteration device time of ``fn`` with events on the current stream."""
    device = torch.device(device if device is not None else "cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(device)
    out = []
    for _ in range(iters):
        if flush:
            flush_l2(device)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize(device)
        out.append(a.elapsed_time(b))
    return out


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while a region is timed."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_ms: int = 200):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self._proc: Optional[subprocess.Popen] = None
        self._lines: List[str] = []
        self._thread: Optional[threading.Thread] = None

    def __enter__(self):
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={self.QUERY}",
                 "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self._thread = threading.Thread(target=self._pump, daemon=True)
            self._thread.start()
        except OSError:
            self._proc = None
        return self

    def _pump(self):
        for line in self._proc.stdout:
            self._lines.append(line.strip())

    def __exit__(self, *exc):
        if self._proc is not None and not self._lines:
            import time
            end = time.time() + 3.0  # short regions: wait for at least one sample
            while not self._lines and time.time() < end:
                time.sleep(0.05)
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
        return False

    def summary(self) -> dict:
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self._lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx),
                "reasons": sorted(reasons), "samples": len(sm)}
