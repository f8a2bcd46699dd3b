"""K9 — alias-method negative sampler (unigram^0.75 in classic word2vec)."""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib


def build_alias(weights):
    """Walker/Vose alias table. Returns (prob float32[V], alias uint32[V])."""
    w = np.asarray(weights, dtype=np.float64)
    v = w.size
    p = w / w.sum() * v
    prob = np.zeros(v, dtype=np.float64)
    alias = np.arange(v, dtype=np.uint32)
    small = np.flatnonzero(p < 1.0).tolist()
    large = np.flatnonzero(p >= 1.0).tolist()
    p = p.copy()
    while small and large:
        s = small.pop()
        l = large.pop()
        prob[s] = p[s]
        alias[s] = l
        p[l] = p[l] + p[s] - 1.0
        (small if p[l] < 1.0 else large).append(l)
    for i in large:
        prob[i] = 1.0
    for i in small:
        prob[i] = 1.0
    return prob.astype(np.float32), alias


class AliasSampler:
    def __init__(self, weights, device, seed=1234):
        prob, alias = build_alias(weights)
        self.vocab = int(prob.size)
        self.prob = torch.from_numpy(prob).to(device)
        self.alias = torch.from_numpy(alias.view(np.int32)).to(device)
        self.seed = int(seed)
        self.offset = 0
        self.device = device

    def sample(self, n: int, out=None, key_offset: int = 0):
        if out is None:
            out = torch.empty(n, dtype=torch.int32, device=self.device)
        _lib.call("ssb_neg_sample", self.prob.data_ptr(), self.alias.data_ptr(), self.vocab, self.seed,
                  self.offset, key_offset, out.data_ptr(), n, _lib.stream_ptr())
        self.offset += (n + 1) // 2
        return out
