"""Small helpers shared by the op wrappers."""
from __future__ import annotations

import numpy as np
import torch


def u32_tensor(x, device=None):
    """Any integer array-like -> torch.int32 tensor holding the uint32 bit pattern."""
    if isinstance(x, torch.Tensor):
        t = x
        if t.dtype == torch.int64:
            t = (t & 0xFFFFFFFF).to(torch.int64)
            t = torch.where(t >= 2 ** 31, t - 2 ** 32, t).to(torch.int32)
        elif t.dtype != torch.int32:
            t = t.to(torch.int32)
        return t.to(device) if device is not None else t
    a = np.ascontiguousarray(np.asarray(x).astype(np.uint32))
    t = torch.from_numpy(a.view(np.int32))
    return t.to(device) if device is not None else t


def to_u32_numpy(t):
    """torch.int32 tensor (uint32 bit pattern) -> numpy uint32."""
    return t.detach().cpu().contiguous().numpy().view(np.uint32)


def check_cuda(*tensors):
    for t in tensors:
        if t is not None and (not t.is_cuda or not t.is_contiguous()):
            raise ValueError("expected contiguous CUDA tensors")
