"""Sparse logistic regression on the parameter server (BASELINE config 3: hashed 2^28 features,
Criteo-shaped synthetic batches, AdaGrad server update).

The reference's ``src/apps/logistic_regression`` is absent from its snapshot; the
loop shape is the framework's (SwiftWorker.h:88-113): collect the minibatch's feature keys → pull →
per-example dot + sigmoid + log-loss, per-feature gradients → push (AdaGrad applied on the owner
shard).  All of it on the GPU: ``ssb_dedup`` → ``ssb_pull`` (dim 1) → ``ssb_lr_fwd_bwd`` →
``ssb_push_adagrad``.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from ..ops import lr as lr_ops
from ..ops.dedup import Deduper
from ..ops.table import ShardedTable


@dataclass
class LRConfig:
    feature_bits: int = 28          # hashed feature space 2^28
    n_numeric: int = 13             # Criteo: 13 integer features + 26 categorical
    n_categorical: int = 26
    batch_size: int = 65536
    lr: float = 0.05
    adagrad_eps: float = 1e-8
    seed: int = 1
    optimizer: str = "adagrad"      # update rule by name (include/ssb200_rule.cuh): sgd | adagrad | ftrl
    l1: float = 0.0                 # FTRL only
    l2: float = 0.0
    fused: bool = True              # one cooperative kernel per step (K6 fused) instead of dedup -> pull -> fwd/bwd -> push
    feature_major: bool = False     # step() is given keys / vals as [n_feat, B] (coalesced loads in the fused kernel)

    @property
    def n_feat(self):
        return self.n_numeric + self.n_categorical


def criteo_like_batch(cfg: LRConfig, n_examples: int, seed: int = 0):
    """Synthetic Criteo-shaped data: keys uint32 [n, 39] hashed into 2^feature_bits, values float32
    [n, 39] (numeric: log1p of a heavy-tailed count; categorical: 1), labels float32 [n] from a
    hidden sparse linear model so there is signal to learn."""
    rng = np.random.default_rng(seed)
    space = 1 << cfg.feature_bits
    keys = np.empty((n_examples, cfg.n_feat), dtype=np.uint64)
    vals = np.ones((n_examples, cfg.n_feat), dtype=np.float32)
    for f in range(cfg.n_numeric):
        keys[:, f] = f  # one weight per numeric field
        vals[:, f] = np.log1p(rng.pareto(1.5, n_examples)).astype(np.float32)
    card = np.array([10 ** (2 + (f % 5)) for f in range(cfg.n_categorical)], dtype=np.int64)  # 1e2 .. 1e6 per field
    for f in range(cfg.n_categorical):
        # Zipf-ish category ids per field
        cat = (card[f] * rng.random(n_examples) ** 3).astype(np.uint64)
        keys[:, cfg.n_numeric + f] = ((np.uint64(f + 1) << np.uint64(40)) + cat)
    # hash into the feature space (fmix64), keep numeric fields at fixed small ids
    from ..parallel.hashfrag import fmix64

    hashed = (fmix64(keys) % np.uint64(space - 64)).astype(np.uint32) + np.uint32(64)
    hashed[:, :cfg.n_numeric] = keys[:, :cfg.n_numeric].astype(np.uint32)
    w_hidden = (fmix64(hashed.astype(np.uint64) ^ np.uint64(0xABCDEF)) % np.uint64(2001)).astype(np.float32) / 1000.0 - 1.0
    z = (w_hidden * vals).sum(1) * 0.3
    labels = (rng.random(n_examples) < 1.0 / (1.0 + np.exp(-z))).astype(np.float32)
    return hashed, vals, labels


class LRWorker:
    def __init__(self, cfg: LRConfig, table: ShardedTable, device):
        assert table.dim == 1 and (table.has_state or cfg.optimizer == "sgd")
        self.cfg, self.table, self.device = cfg, table, device
        self.loss = torch.zeros(2, dtype=torch.float32, device=device)
        self.fused = None
        if cfg.fused:
            self.fused = lr_ops.FusedLRState(table, cfg.batch_size, cfg.n_feat, cfg.feature_bits, device)
            self.n_unique = self.fused.counters[0:1]
            return
        n = cfg.batch_size * cfg.n_feat
        i32 = dict(dtype=torch.int32, device=device)
        self.unique = torch.empty(n, **i32)
        self.inverse = torch.empty(n, **i32)
        self.n_unique = torch.zeros(1, **i32)
        self.gslot = torch.empty(n, **i32)
        self.params = torch.empty((n, 1), dtype=torch.float32, device=device)
        self.grads = torch.zeros((n, 1), dtype=torch.float32, device=device)
        self.dedup = Deduper(n, device)
        self.n = n

    def step(self, keys: torch.Tensor, vals: torch.Tensor, labels: torch.Tensor):
        """keys int32 [B, n_feat] (uint32 bits), vals float32 [B, n_feat], labels float32 [B]."""
        cfg = self.cfg
        if self.fused is not None:
            hy = dict(lr=cfg.lr, eps=cfg.adagrad_eps)
            if cfg.optimizer == "ftrl":
                hy = dict(lr=cfg.lr, eps=1.0, a=cfg.l1, b=cfg.l2)
            lr_ops.lr_fused_step(self.fused, keys, vals, labels, self.loss, rule=cfg.optimizer, feature_major=cfg.feature_major, **hy)
            return self.loss
        if cfg.optimizer != "adagrad":
            raise ValueError("the unfused path implements AdaGrad only; use fused=True for %r" % cfg.optimizer)
        self.dedup(keys.reshape(-1), unique=self.unique, inverse=self.inverse, n_unique=self.n_unique)
        self.table.pull(self.unique, n_dev=self.n_unique, rows_out=self.params, gslot_out=self.gslot,
                        grads_zero=self.grads, n=self.n)
        lr_ops.lr_fwd_bwd(self.params, self.grads, self.inverse, vals, labels, cfg.n_feat, self.loss)
        self.table.push_adagrad(self.gslot, self.grads, cfg.lr, cfg.adagrad_eps, n_dev=self.n_unique,
                                zero_grads=False, n=self.n)
        return self.loss

    def reset_loss(self):
        self.loss.zero_()

    def health(self):
        """Counters that must stay zero (read on the host): touched-list overflow, keys that found no slot."""
        if self.fused is None:
            return {"table_full": int(self.table.stats[1].item())}
        c = self.fused.counters.tolist()
        return {"touched": c[0], "touched_overflow": c[2], "table_full": c[3] + int(self.table.stats[1].item())}

    def mean_loss(self):
        s, n = self.loss.tolist()
        return s / max(n, 1.0)
