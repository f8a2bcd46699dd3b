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
    tile_impl: str = "cuda_core"  # "cuda_core" | "tcgen05" | "tcgen05_direct" (fused pull: TMA gathers from the table's bf16 mirror)
    # worker-side key -> slot cache: repeat pulls skip the owner's key-bucket probe. None = auto: on when some
    # owners are remote (saves an NVLink round trip per key: 0.769 -> 0.712 ms/step at 2 GPUs), off on one GPU
    # (the bucket probe is an L2 hit there and the extra look-up costs 10% of the pull)
    slot_cache: bool | None = None


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


class _StepBuffers:
    """One set of per-step device buffers (the worker cache for one minibatch)."""

    def __init__(self, n_keys, dim, max_pairs, device, tc):
        i32 = dict(dtype=torch.int32, device=device)
        self.keys_all = torch.empty(n_keys, **i32)
        self.unique = torch.empty(n_keys, **i32)
        self.inverse = torch.empty(n_keys, **i32)
        self.n_unique = torch.zeros(1, **i32)
        self.gslot = torch.empty(n_keys, **i32)
        self.params = None if tc else torch.empty((n_keys, dim), dtype=torch.float32, device=device)
        self.params_bf16 = torch.empty((n_keys, dim), dtype=torch.bfloat16, device=device) if tc else None
        self.grads = torch.zeros((n_keys, dim), dtype=torch.float32, device=device)
        self.pairs_in = torch.empty(max(max_pairs, 1), **i32)
        self.pairs_out = torch.empty(max(max_pairs, 1), **i32)
        self.n_pairs = torch.zeros(1, **i32)
        self.dedup = Deduper(n_keys, device)
        self.pulled = torch.cuda.Event()
        self.computed = torch.cuda.Event()
        self.freed = torch.cuda.Event()
        self.step_idx = 0


class Word2VecWorker:
    """The worker-side algorithm (``BaseAlgorithm`` analogue) for one GPU.

    ``pipeline=True`` runs the communication half of a step (key build, negative sampling, dedup,
    pull — the part that crosses NVLink) on its own stream, one step ahead of the compute half
    (SGNS tiles + push), with two buffer sets.  A pull may then read a row before the previous
    step's push has landed: bounded staleness of one step, well inside what the reference's
    fully asynchronous workers already allow (no cross-worker synchronisation)."""

    def __init__(self, cfg: Word2VecConfig, table: ShardedTable, device, sampler: AliasSampler | None = None,
                 pipeline: bool = False):
        assert table.dim == cfg.dim
        self.cfg, self.table, self.device = cfg, table, device
        B, w = cfg.tokens_per_step, cfg.window
        self.tc = cfg.tile_impl in ("tcgen05", "tcgen05_direct")
        self.direct = cfg.tile_impl == "tcgen05_direct"
        if self.direct and getattr(table, "mirror_map", None) is None:
            raise ValueError("tcgen05_direct needs a table with a single-range bf16 row mirror (with_shadow=True)")
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
            # synthetic-benchmark default: the noise distribution of the synthetic Zipf corpus. Training from a real
            # corpus replaces it with the corpus' own unigram^0.75 (Word2VecTrainer.train_file). The rank is mixed
            # into the seed so that the workers of a job do not all draw the same negatives.
            sampler = AliasSampler(unigram_weights_from_zipf(cfg.vocab, cfg.zipf_s), device,
                                   seed=cfg.seed + 77 + 7919 * int(getattr(table, "rank", 0)))
        self.sampler = sampler
        use_cache = cfg.slot_cache if cfg.slot_cache is not None else getattr(table, "num_shards", 1) > 1
        if use_cache and getattr(table, "slot_cache", 0) is None:
            table.enable_slot_cache(2 * cfg.vocab, dense=True)     # keys = word ids: input w, output w + vocab
        self.pipeline = bool(pipeline)
        self.has_padding = False       # set when batches may contain padding tokens (ragged sentences)
        self.sets = [_StepBuffers(self.n_keys, cfg.dim, self.max_pairs, device, self.tc)
                     for _ in range(2 if self.pipeline else 1)]
        self.loss = torch.zeros(2, dtype=torch.float32, device=device)
        self.step_idx = 0
        self.neg_weight = float(cfg.negatives) / float(cfg.neg_pool)
        self.tma_mode = 33      # kernel mode bits: 1 = TMA tensor map made (gather4 path available), 32 = gather by cp.async warps (faster: 212 vs 261 us)
        if self.pipeline:
            lo, hi = torch.cuda.Stream.priority_range()
            self.comm_stream = torch.cuda.Stream(device=device, priority=lo)
            self.comp_stream = torch.cuda.Stream(device=device, priority=hi)
            self.push_stream = torch.cuda.Stream(device=device, priority=lo)
            # optional cap on the grid of the grid-stride comm kernels (blocks per SM). Measured: no
            # cap is best (0.362 ms/step vs 0.366 at 5, 0.380 at 3, 0.388 at 2 on one B200)
            import os

            _lib.kernels().ssb_set_blocks_per_sm_cap(int(os.environ.get("SSB200_BLOCKS_PER_SM_CAP", "0")))
        # aliases for the non-pipelined path / introspection
        self.n_unique = self.sets[0].n_unique
        self.grads = self.sets[0].grads
        self.params = self.sets[0].params
        self.params_bf16 = self.sets[0].params_bf16

    # ---- the two halves of a step ----------------------------------------------------------
    def _comm(self, s, tokens):
        cfg, B = self.cfg, self.cfg.tokens_per_step
        ka = s.keys_all
        # input keys = w, output keys = w + vocab (padding tokens stay padding): one kernel
        _lib.call("ssb_build_keys", tokens.data_ptr(), B, cfg.vocab, ka.data_ptr(), _lib.stream_ptr())
        self.sampler.sample(self.n_negs, out=ka[2 * B:], key_offset=cfg.vocab)
        # K2: unique keys + inverse index (the cache has one entry per key)
        s.dedup(ka, unique=s.unique, inverse=s.inverse, n_unique=s.n_unique)
        # K1+K3+K4: route, probe/insert, gather into the cache; grads zeroed in the same pass
        self.table.pull(s.unique, n_dev=s.n_unique, rows_out=s.params, rows_bf16=None if self.direct else s.params_bf16,
                        want_fp32=not self.tc, gslot_out=s.gslot, grads_zero=s.grads, n=self.n_keys)

    def _compute(self, s):
        cfg, B = self.cfg, self.cfg.tokens_per_step
        if self.direct:
            # fused pull: the pull above only resolved slots and reset the gradients; the operand rows
            # are gathered by TMA straight from the owners' bf16 row mirrors (own HBM or NVLink peers)
            sgns_ops.sgns_block_tc_direct(self.table, s.gslot, s.grads, s.inverse[:B], s.inverse[B:2 * B],
                                          cfg.sent_len, cfg.window, cfg.seed, s.step_idx * B, s.inverse[2 * B:],
                                          float(cfg.negatives), self.loss, mode=self.tma_mode)
        elif self.tc:
            # K5 on tensor cores: TMA gather4 -> tcgen05.mma (bf16) -> TMEM epilogue -> red
            sgns_ops.sgns_block_tc(s.params_bf16, s.grads, s.inverse[:B], s.inverse[B:2 * B],
                                   cfg.sent_len, cfg.window, cfg.seed, s.step_idx * B, s.inverse[2 * B:],
                                   float(cfg.negatives), self.loss, mode=self.tma_mode)
        else:
            # pairs over cache row indices, then the CUDA-core tile kernel
            sgns_ops.make_pairs(s.inverse[:B], s.inverse[B:2 * B], cfg.sent_len, cfg.window,
                                cfg.seed, s.step_idx * B, s.pairs_in, s.pairs_out, s.n_pairs)
            sgns_ops.sgns_tiles(s.params, s.grads, s.pairs_in, s.pairs_out, s.n_pairs,
                                self.max_pairs, s.inverse[2 * B:], cfg.neg_pool, self.neg_weight, self.loss)

    def _push(self, s):
        cfg = self.cfg
        # K7/K8: scatter-add fused with the server update
        inbox = getattr(self.table, "inbox", None)
        if inbox is not None and cfg.optimizer == "adagrad":
            # multi-GPU AdaGrad: ship gradients once, the owner applies (one NVLink crossing per row)
            inbox.push(s.gslot, s.grads, cfg.lr, cfg.adagrad_eps, adagrad=True, n_dev=s.n_unique, n=self.n_keys)
        elif cfg.optimizer == "adagrad":
            self.table.push_adagrad(s.gslot, s.grads, cfg.lr, cfg.adagrad_eps, n_dev=s.n_unique,
                                    zero_grads=False, n=self.n_keys)
        else:
            self.table.push_sgd(s.gslot, s.grads, cfg.lr, n_dev=s.n_unique, zero_grads=False, n=self.n_keys)

    def step(self, tokens: torch.Tensor):
        """One minibatch: tokens int32 [tokens_per_step] on the device. Returns the loss tensor
        [sum_loss, n_pairs] (device, accumulated since the last ``reset_loss``)."""
        assert tokens.numel() == self.cfg.tokens_per_step
        s = self.sets[self.step_idx % len(self.sets)]
        s.step_idx = self.step_idx
        if not self.pipeline:
            self._comm(s, tokens)
            self._compute(s)
            self._push(s)
        else:
            # keep the host at most two steps ahead of the device: deeper queues on the two
            # streams measurably reduce their overlap (the e2e arm, which is throttled by its loss
            # read-back, was 20% faster than the free-running device arm before this)
            s.freed.synchronize()
            cur = torch.cuda.current_stream()
            self.comm_stream.wait_stream(cur)              # tokens were produced on the caller's stream
            with torch.cuda.stream(self.comm_stream):
                self.comm_stream.wait_event(s.freed)        # this buffer set's previous push has finished
                self._comm(s, tokens)
                s.pulled.record(self.comm_stream)
            with torch.cuda.stream(self.comp_stream):
                self.comp_stream.wait_event(s.pulled)
                self._compute(s)
                s.computed.record(self.comp_stream)
            with torch.cuda.stream(self.push_stream):       # push(i) overlaps the SGNS tiles of step i+1
                self.push_stream.wait_event(s.computed)
                self._push(s)
                s.freed.record(self.push_stream)
            tokens.record_stream(self.comm_stream)
        self.step_idx += 1
        return self.loss

    def flush(self):
        """Make the caller's stream wait for everything enqueued so far (pipelined mode)."""
        if self.pipeline:
            cur = torch.cuda.current_stream()
            cur.wait_stream(self.comm_stream)
            cur.wait_stream(self.comp_stream)
            cur.wait_stream(self.push_stream)

    def reset_loss(self):
        if self.pipeline:
            with torch.cuda.stream(self.comp_stream):
                self.loss.zero_()
        else:
            self.loss.zero_()

    def mean_loss(self):
        self.flush()
        s, n = self.loss.tolist()
        return s / max(n, 1.0)


class Word2VecTrainer:
    """Public end-to-end API: feed host token batches, get losses back.

    ``train_step(tokens_host)`` copies the batch host->device from pinned memory, runs the
    worker step, and issues an asynchronous device->host copy of the step's loss into a pinned
    slot; it returns the *previous* step's mean loss (None on the first call) so the host never
    stalls the GPU.  ``finish()`` returns the last loss.  Bytes per step: ``h2d_bytes`` /
    ``d2h_bytes``."""

    def __init__(self, cfg: Word2VecConfig, table: ShardedTable, device, pipeline: bool = False):
        self.worker = Word2VecWorker(cfg, table, device, pipeline=pipeline)
        self.cfg, self.device = cfg, device
        self.tok_dev = [torch.empty(cfg.tokens_per_step, dtype=torch.int32, device=device) for _ in range(2)]
        self.loss_host = torch.zeros((2, 2), dtype=torch.float32).pin_memory()
        self.events = [torch.cuda.Event(), torch.cuda.Event()]
        self.h2d_bytes = cfg.tokens_per_step * 4
        self.d2h_bytes = 2 * 4
        self._i = 0
        self.on_step = None        # optional callback(step) after every minibatch of train_file (periodic backups)
        self.bad_tokens = 0
        from ..utils.metrics import registry

        reg = registry()
        self._m_tokens = reg.counter("ssb_w2v_tokens_total", "tokens fed to Word2VecTrainer.train_step")
        self._m_steps = reg.counter("ssb_w2v_steps_total", "minibatches trained")
        self._m_rate = reg.meter("ssb_w2v_words_per_second", "tokens per second over the last 5 s")
        self._m_loss = reg.gauge("ssb_w2v_loss_per_pair", "mean SGNS loss of the latest completed step")

    def train_step(self, tokens_host: torch.Tensor):
        i = self._i & 1
        self.tok_dev[i].copy_(tokens_host, non_blocking=True)
        w = self.worker
        if w.pipeline:
            # per-step loss: snapshot-and-clear on the compute stream right after the step's kernels
            w.step(self.tok_dev[i])
            with torch.cuda.stream(w.comp_stream):
                self.loss_host[i].copy_(w.loss, non_blocking=True)
                w.loss.zero_()
                self.events[i].record(w.comp_stream)
        else:
            w.reset_loss()
            loss = w.step(self.tok_dev[i])
            self.loss_host[i].copy_(loss, non_blocking=True)
            self.events[i].record()
        prev = None
        if self._i > 0:
            j = 1 - i
            self.events[j].synchronize()
            s, n = self.loss_host[j].tolist()
            prev = s / max(n, 1.0)
            self._m_loss.set(prev)
        self._i += 1
        self._m_tokens.inc(self.cfg.tokens_per_step)
        self._m_steps.inc()
        self._m_rate.mark(self.cfg.tokens_per_step)
        return prev

    def health(self):
        """Counters of the two ways an update can be lost, read from the device (synchronises): rows the table had
        no room for (``table_full``: their pull returned zeros and their push was skipped) and gradient rows dropped
        because a sender's share of an owner's inbox was full (``inbox_overflow``). Both must stay 0."""
        self.worker.flush()
        out = {"table_full": int(self.worker.table.stats[1].item()), "inbox_overflow": 0}
        inbox = getattr(self.worker.table, "inbox", None)
        if inbox is not None:
            out["inbox_overflow"] = int(inbox.overflow.item())
            if hasattr(inbox, "err"):
                out["inbox_overflow"] += int(inbox.err.item())      # an apply wait ran out: rows were skipped
        return out

    def check_health(self):
        h = self.health()
        if h["table_full"] or h["inbox_overflow"]:
            raise RuntimeError("updates were lost: %r (grow the table capacity / the inbox)" % (h,))
        return h

    def use_corpus_noise(self, path, power=0.75):
        """Negative-sampling table from the corpus' own word counts (unigram^power), as word2vec defines it.
        Returns (distinct words seen, ids >= vocab found in the corpus)."""
        from ..utils.corpus import count_words

        counts, bad = count_words(path, self.cfg.vocab)
        seen = int((counts > 0).sum())
        if seen > 0:
            w = self.worker
            w.sampler = AliasSampler(np.power(counts.astype(np.float64), power), self.device,
                                     seed=self.cfg.seed + 77 + 7919 * int(getattr(w.table, "rank", 0)))
        return seen, bad

    def train_file(self, path, epochs=1, log_every=0, split=None, max_steps=None, corpus_noise=True):
        """Train from a text corpus (one sentence of integer word ids per line — the reference's data
        format, tools/gen_word2vec_data.py). Sentences are cut / padded to ``cfg.sent_len``; a native
        parallel reader (utils/corpus.py) fills pinned batches while the GPU works.
        ``split=(n, i)`` reads worker i's share of the lines; ``max_steps`` stops after that many
        minibatches (with the inbox push every rank has to run the same number of steps).
        ``corpus_noise``: draw negatives from the corpus' own unigram^0.75 (counted here, before training).
        Token ids >= vocab are dropped (turned into padding) and counted in ``self.bad_tokens``; lost updates
        (table full / inbox overflow) raise. Returns (tokens_trained, last_mean_loss)."""
        from ..utils.corpus import CorpusLoader

        if corpus_noise:
            self.use_corpus_noise(path)
        self.worker.has_padding = True
        loader = CorpusLoader(path, self.cfg.tokens_per_step, self.cfg.sent_len, epochs=epochs, split=split,
                              max_id=self.cfg.vocab)
        total, step = 0, 0
        for buf, n_real in loader:
            if max_steps is not None and step >= max_steps:
                break
            loss = self.train_step(buf)
            total += n_real
            step += 1
            if self.on_step is not None:
                self.on_step(step)
            if log_every and step % log_every == 0 and loss is not None:
                self.check_health()
                print("step %d  tokens %d  loss/pair %.4f" % (step, total, loss))
        self.bad_tokens = loader.bad_tokens
        loader.close()
        self.steps_done = step
        last = self.finish()
        self.check_health()
        return total, last

    def finish(self):
        if self._i == 0:
            return None
        j = (self._i - 1) & 1
        self.events[j].synchronize()
        s, n = self.loss_host[j].tolist()
        return s / max(n, 1.0)
