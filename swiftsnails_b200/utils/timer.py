"""Timing / tracing helpers (the reference has only a seconds-resolution wall
timer, utils/Timer.h:14-44).  ``Timer`` is the wall-clock twin of the C++ one; ``CudaTimer``
times a region with CUDA events on the current stream (the rule for every GPU number in this
repo); ``nvtx_range`` annotates phases for Nsight tools."""
from __future__ import annotations

import contextlib
import time


class Timer:
    def __init__(self):
        self.start()

    def start(self):
        self._t0 = time.perf_counter()
        self._t1 = None

    def stop(self):
        self._t1 = time.perf_counter()
        return self.elapsed()

    def elapsed(self):
        return (self._t1 if self._t1 is not None else time.perf_counter()) - self._t0


class CudaTimer:
    """with CudaTimer() as t: ...  ->  t.ms (after the region, synchronises the end event)."""

    def __init__(self, stream=None):
        import torch

        self._torch, self._stream = torch, stream
        self.ms = None

    def __enter__(self):
        torch = self._torch
        self._e0, self._e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self._e0.record(self._stream)
        return self

    def __exit__(self, *exc):
        self._e1.record(self._stream)
        self._e1.synchronize()
        self.ms = self._e0.elapsed_time(self._e1)
        return False


def max_over_ranks(ms, device):
    """Every multi-GPU number is the max over ranks of a device-side time."""
    import torch
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size() == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


@contextlib.contextmanager
def nvtx_range(name):
    import torch

    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()
