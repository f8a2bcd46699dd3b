"""Large embedding table pull/push worker (BASELINE configs 4 and 5): random key batches against a
table sharded across the GPUs, AdaGrad or SGD on push.  Used by ``benchmarks/pull_push_sweep.py``
and ``benchmarks/capacity_200m.py``.  The reference analogue is the synthetic pull/push loop run in
SURVEY.md Appendix A (pull_with_barrier / push_with_barrier on unique uint32 keys)."""
from __future__ import annotations

import torch

from ..ops.dedup import Deduper


class PullPushWorker:
    def __init__(self, table, max_keys, device, optimizer="sgd", lr=0.01, bf16_cache=False):
        self.table, self.device, self.optimizer, self.lr = table, device, optimizer, lr
        self.max_keys = int(max_keys)
        self.gslot = torch.empty(self.max_keys, dtype=torch.int32, device=device)
        self.rows = torch.empty((self.max_keys, table.dim), dtype=torch.float32, device=device)
        self.grads = torch.full((self.max_keys, table.dim), 1e-3, dtype=torch.float32, device=device)
        self.unique = torch.empty(self.max_keys, dtype=torch.int32, device=device)
        self.inverse = torch.empty(self.max_keys, dtype=torch.int32, device=device)
        self.n_unique = torch.zeros(1, dtype=torch.int32, device=device)
        self.dedup = Deduper(self.max_keys, device)

    def pull(self, keys):
        n = keys.numel()
        self.table.pull(keys, rows_out=self.rows, gslot_out=self.gslot, n=n)
        return self.rows[:n]

    def push(self, n):
        if self.optimizer == "adagrad":
            self.table.push_adagrad(self.gslot, self.grads, self.lr, n=n, zero_grads=False)
        else:
            self.table.push_sgd(self.gslot, self.grads, self.lr, n=n, zero_grads=False)

    def dedup_pull_push(self, keys):
        """Full worker step on a batch with duplicates: dedup → pull unique → push unique."""
        self.dedup(keys, unique=self.unique, inverse=self.inverse, n_unique=self.n_unique)
        self.table.pull(self.unique, n_dev=self.n_unique, rows_out=self.rows, gslot_out=self.gslot, n=keys.numel())
        if self.optimizer == "adagrad":
            self.table.push_adagrad(self.gslot, self.grads, self.lr, n_dev=self.n_unique, n=keys.numel(), zero_grads=False)
        else:
            self.table.push_sgd(self.gslot, self.grads, self.lr, n_dev=self.n_unique, n=keys.numel(), zero_grads=False)
