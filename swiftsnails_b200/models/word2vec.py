"""word2vec skip-gram with negative sampling on the parameter server (BASELINE config 2).

The reference's ``src/apps/word2vec`` is absent from its snapshot; the shape of
the worker loop is dictated by the framework
(/root/reference/src/core/framework/SwiftWorker.h:19-57,88-113 and SURVEY.md §3.4):

    collect minibatch keys -> pull_with_barrier(keys, cache) -> dot products + sigmoid,
    accumulate cache.grads()[k] -> push_with_barrier(keys, cache)

Here the whole loop body runs on the GPU: keys are de-duplicated by ``ssb_dedup`` (the
cache's one-entry-per-key property), ``ssb_pull`` gathers the rows from their owner shards
(local HBM or a peer over NVLink) straight into the dense worker cache, the SGNS tile kernel
accumulates gradients into the cache's grad buffer, and ``ssb_push_*`` scatter-adds the
update into the owner rows.  Keys: input vector of word ``w`` is key ``w``, its output
(context) vector is key ``w + vocab`` — one flat uint32 key space, as the reference requires
(Key = index_t).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from .. import _lib
from ..ops import sgns as sgns_ops
from ..ops.dedup import Deduper
from ..ops.sampling import AliasSampler
from ..ops.table import ShardedTable


@dataclass
class Word2VecConfig:
    vocab: int = 1_000_000
    dim: int = 128
    window: int = 5
    negatives: int = 5          # K of the objective (per pair, in expectation)
    neg_pool: int = 16          # negatives actually drawn per 128-pair tile, shared by the tile
    sent_len: int = 1000        # tokens per synthetic sentence (pairs never cross sentences)
    tokens_per_step: int = 65536
    lr: float = 0.025
    optimizer: str = "sgd"      # "sgd" | "adagrad"
    adagrad_eps: float = 1e-8
    zipf_s: float = 1.0
    seed: int = 1
    tile_impl: str = "cuda_core"  # "cuda_core" | "tcgen05"


def zipf_weights(vocab: int, s: float = 1.0):
    return 1.0 / np.power(np.arange(1, vocab + 1, dtype=np.float64), s)


def synthetic_corpus(vocab: int, n_tokens: int, s: float = 1.0, seed: int = 0, permute=True):
    """Zipf-distributed token stream (uint32).  Word ids are a fixed permutation of the ranks so
    frequency is not correlated with id."""
    w = zipf_weights(vocab, s)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    rng = np.random.default_rng(seed)
    ranks = np.searchsorted(cdf, rng.random(n_tokens), side="right").astype(np.int64)
    ranks = np.minimum(ranks, vocab - 1)
    if permute:
        perm = np.random.default_rng(12345).permutation(vocab)
        ranks = perm[ranks]
    return ranks.astype(np.uint32)


def unigram_weights_from_zipf(vocab: int, s: float = 1.0, power: float = 0.75, permute=True):
    w = np.power(zipf_weights(vocab, s), power)
    if permute:
        perm = np.random.default_rng(12345).permutation(vocab)
        out = np.empty_like(w)
        out[perm] = w
        return out
    return w


class Word2VecWorker:
    """The worker-side algorithm (``BaseAlgorithm`` analogue) for one GPU."""

    def __init__(self, cfg: Word2VecConfig, table: ShardedTable, device, sampler: AliasSampler | None = None):
        assert table.dim == cfg.dim
        self.cfg, self.table, self.device = cfg, table, device
        B, w = cfg.tokens_per_step, cfg.window
        self.tc = cfg.tile_impl == "tcgen05"
        if self.tc:
            # block formulation (csrc/kernels/sgns_tc.cu): a tile = 128 consecutive centre tokens
            # sharing 16 negatives; no pair list is materialised
            assert cfg.dim == 128 and cfg.window <= 8 and cfg.neg_pool == 16
            self.max_pairs = 0
            self.max_tiles = (B + 127) // 128
        else:
            self.max_pairs = B * 2 * w
            self.max_tiles = (self.max_pairs + 127) // 128
        self.n_negs = self.max_tiles * cfg.neg_pool
        self.n_keys = 2 * B + self.n_negs
        if sampler is None:
            sampler = AliasSampler(unigram_weights_from_zipf(cfg.vocab, cfg.zipf_s), device, seed=cfg.seed + 77)
        self.sampler = sampler
        i32 = dict(dtype=torch.int32, device=device)
        self.keys_all = torch.empty(self.n_keys, **i32)
        self.unique = torch.empty(self.n_keys, **i32)
        self.inverse = torch.empty(self.n_keys, **i32)
        self.n_unique = torch.zeros(1, **i32)
        self.gslot = torch.empty(self.n_keys, **i32)
        self.params = None if self.tc else torch.empty((self.n_keys, cfg.dim), dtype=torch.float32, device=device)
        self.params_bf16 = torch.empty((self.n_keys, cfg.dim), dtype=torch.bfloat16, device=device) if self.tc else None
        self.grads = torch.zeros((self.n_keys, cfg.dim), dtype=torch.float32, device=device)
        self.pairs_in = torch.empty(max(self.max_pairs, 1), **i32)
        self.pairs_out = torch.empty(max(self.max_pairs, 1), **i32)
        self.tma_mode = 1
        self.n_pairs = torch.zeros(1, **i32)
        self.loss = torch.zeros(2, dtype=torch.float32, device=device)
        self.dedup = Deduper(self.n_keys, device)
        self.step_idx = 0
        self.neg_weight = float(cfg.negatives) / float(cfg.neg_pool)

    def step(self, tokens: torch.Tensor):
        """One minibatch: tokens int32 [tokens_per_step] on the device. Returns the loss tensor
        [sum_loss, n_pairs] (device, accumulated since the last ``reset_loss``)."""
        cfg, B = self.cfg, self.cfg.tokens_per_step
        assert tokens.numel() == B
        ka = self.keys_all
        ka[:B].copy_(tokens)                               # input keys  = w
        torch.add(tokens, cfg.vocab, out=ka[B:2 * B])      # output keys = w + vocab
        self.sampler.sample(self.n_negs, out=ka[2 * B:], key_offset=cfg.vocab)
        # K2: unique keys + inverse index (the cache has one entry per key)
        self.dedup(ka, unique=self.unique, inverse=self.inverse, n_unique=self.n_unique)
        # K1+K3+K4: route, probe/insert, gather into the cache; grads zeroed in the same pass
        self.table.pull(self.unique, n_dev=self.n_unique, rows_out=self.params, rows_bf16=self.params_bf16,
                        want_fp32=not self.tc, gslot_out=self.gslot, grads_zero=self.grads, n=self.n_keys)
        if self.tc:
            # K5 on tensor cores: TMA gather4 -> tcgen05.mma (bf16) -> TMEM epilogue -> red
            sgns_ops.sgns_block_tc(self.params_bf16, self.grads, self.inverse[:B], self.inverse[B:2 * B],
                                   cfg.sent_len, cfg.window, cfg.seed, self.step_idx * B, self.inverse[2 * B:],
                                   float(cfg.negatives), self.loss, mode=self.tma_mode)
        else:
            # pairs over cache row indices, then the CUDA-core tile kernel
            sgns_ops.make_pairs(self.inverse[:B], self.inverse[B:2 * B], cfg.sent_len, cfg.window,
                                cfg.seed, self.step_idx * B, self.pairs_in, self.pairs_out, self.n_pairs)
            sgns_ops.sgns_tiles(self.params, self.grads, self.pairs_in, self.pairs_out, self.n_pairs,
                                self.max_pairs, self.inverse[2 * B:], cfg.neg_pool, self.neg_weight, self.loss)
        # K7/K8: scatter-add fused with the server update
        if cfg.optimizer == "adagrad":
            self.table.push_adagrad(self.gslot, self.grads, cfg.lr, cfg.adagrad_eps, n_dev=self.n_unique,
                                    zero_grads=False, n=self.n_keys)
        else:
            self.table.push_sgd(self.gslot, self.grads, cfg.lr, n_dev=self.n_unique, zero_grads=False,
                                n=self.n_keys)
        self.step_idx += 1
        return self.loss

    def reset_loss(self):
        self.loss.zero_()

    def mean_loss(self):
        s, n = self.loss.tolist()
        return s / max(n, 1.0)


class Word2VecTrainer:
    """Public end-to-end API: feed host token batches, get losses back.

    ``train_step(tokens_host)`` copies the batch host->device from pinned memory, runs the
    worker step, and issues an asynchronous device->host copy of the step's loss into a pinned
    slot; it returns the *previous* step's mean loss (None on the first call) so the host never
    stalls the GPU.  ``finish()`` returns the last loss.  Bytes per step: ``h2d_bytes`` /
    ``d2h_bytes``."""

    def __init__(self, cfg: Word2VecConfig, table: ShardedTable, device):
        self.worker = Word2VecWorker(cfg, table, device)
        self.cfg, self.device = cfg, device
        self.tok_dev = [torch.empty(cfg.tokens_per_step, dtype=torch.int32, device=device) for _ in range(2)]
        self.loss_host = torch.zeros((2, 2), dtype=torch.float32).pin_memory()
        self.events = [torch.cuda.Event(), torch.cuda.Event()]
        self.h2d_bytes = cfg.tokens_per_step * 4
        self.d2h_bytes = 2 * 4
        self._i = 0

    def train_step(self, tokens_host: torch.Tensor):
        i = self._i & 1
        self.tok_dev[i].copy_(tokens_host, non_blocking=True)
        self.worker.reset_loss()
        loss = self.worker.step(self.tok_dev[i])
        self.loss_host[i].copy_(loss, non_blocking=True)
        self.events[i].record()
        prev = None
        if self._i > 0:
            j = 1 - i
            self.events[j].synchronize()
            s, n = self.loss_host[j].tolist()
            prev = s / max(n, 1.0)
        self._i += 1
        return prev

    def finish(self):
        if self._i == 0:
            return None
        j = (self._i - 1) & 1
        self.events[j].synchronize()
        s, n = self.loss_host[j].tolist()
        return s / max(n, 1.0)
