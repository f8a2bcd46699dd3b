"""K2 — key de-duplication with inverse index (the GPU form of the reference's
``GlobalParamCache`` maps holding one entry per unique key,
/root/reference/src/core/parameter/global_param_cache.h:107-108)."""
from __future__ import annotations

import torch

from .. import _lib
from ._util import check_cuda


def _pow2_at_least(n):
    p = 64
    while p < n:
        p <<= 1
    return p


class Deduper:
    """Owns the scratch hash set; one instance per stream of batches."""

    def __init__(self, max_keys: int, device):
        self.max_keys = int(max_keys)
        self.cap = _pow2_at_least(2 * self.max_keys)
        self.set = torch.zeros(self.cap, dtype=torch.int64, device=device)
        self.set_ids = torch.empty(self.cap, dtype=torch.int32, device=device)
        self.epoch = 0
        self.device = device

    def __call__(self, keys: torch.Tensor, unique=None, inverse=None, n_unique=None):
        check_cuda(keys)
        n = keys.numel()
        if n > self.max_keys:
            raise ValueError("batch of %d keys exceeds Deduper capacity %d" % (n, self.max_keys))
        if unique is None:
            unique = torch.empty(n, dtype=torch.int32, device=keys.device)
        if inverse is None:
            inverse = torch.empty(n, dtype=torch.int32, device=keys.device)
        if n_unique is None:
            n_unique = torch.empty(1, dtype=torch.int32, device=keys.device)
        self.epoch += 1
        if self.epoch >= 2 ** 32 - 1:
            self.set.zero_()
            self.epoch = 1
        _lib.call("ssb_dedup", keys.data_ptr(), n, self.set.data_ptr(), self.set_ids.data_ptr(), self.cap,
                  self.epoch, unique.data_ptr(), inverse.data_ptr(), n_unique.data_ptr(), 1, _lib.stream_ptr())
        return unique, inverse, n_unique
