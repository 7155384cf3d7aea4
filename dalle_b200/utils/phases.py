"""Phase tracing for the training step: NVTX ranges plus CUDA-event timings per phase.

The reference has no profiler hooks; here every phase of a step (forward, backward, averaging+optimizer) can be
bracketed at no cost when disabled.  Enable with ``DALLE_B200_TRACE=1`` (or ``PhaseTrace(enabled=True)``): NVTX
ranges show up in Nsight tools, and ``summary()`` returns device-timed milliseconds per phase - events are recorded
on the launching stream and only read back in ``summary()``, so tracing does not serialise the step.
"""
from __future__ import annotations

import contextlib
import os
from collections import defaultdict
from typing import Dict, List, Tuple

import torch


class PhaseTrace:
    def __init__(self, enabled: bool | None = None):
        if enabled is None:
            enabled = os.environ.get("DALLE_B200_TRACE", "") not in ("", "0")
        self.enabled = bool(enabled) and torch.cuda.is_available()
        self._pending: List[Tuple[str, torch.cuda.Event, torch.cuda.Event]] = []
        self._totals: Dict[str, List[float]] = defaultdict(list)

    @contextlib.contextmanager
    def phase(self, name: str):
        if not self.enabled:
            yield
            return
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.nvtx.range_push(name)
        start.record()
        try:
            yield
        finally:
            end.record()
            torch.cuda.nvtx.range_pop()
            self._pending.append((name, start, end))

    def summary(self, reset: bool = True) -> Dict[str, float]:
        """Mean device milliseconds per phase since the last call (synchronises once)."""
        if not self.enabled:
            return {}
        torch.cuda.synchronize()
        for name, s, e in self._pending:
            self._totals[name].append(s.elapsed_time(e))
        self._pending.clear()
        out = {k: sum(v) / len(v) for k, v in self._totals.items() if v}
        if reset:
            self._totals.clear()
        return out
