"""The HBM-resident sharded sparse table and its pull / push ops (K3, K4, K7, K8, K11).

ref: /root/reference/src/core/parameter/sparsetable.h:69-198 (SparseTable, PullAccessAgent,
PushAccessAgent), global_pull_access.h:40-120, global_push_access.h:36-149,
system/server/init.h:49-69,116-149.

One *shard* per GPU ("server"): ``keys[cap]`` / ``rows[cap, D]`` / ``state[cap, D]``.  A
``ShardedTable`` is one process's view of all shards — its own plus every peer's mapped over
NVLink (``swiftsnails_b200.parallel.peer``) — and pull/push are single kernels that address the
owner shard directly.  For tests, several shards may live on one GPU (``local_shards=S``): the
kernels do not care where a shard pointer points.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .. import _lib
from ..parallel.hashfrag import HashFrag
from ._util import check_cuda


def round_capacity(rows_expected: int, load_factor: float = 0.5) -> int:
    cap = int(rows_expected / load_factor) + _lib.BUCKET
    return (cap + _lib.BUCKET - 1) // _lib.BUCKET * _lib.BUCKET


class DeviceShard:
    """Storage of one shard on one device."""

    def __init__(self, capacity: int, dim: int, device, with_state, alloc=None, with_shadow: bool = False):
        # with_state: False / True (one word per weight: AdaGrad) or the number of optimizer-state words per weight
        assert capacity % _lib.BUCKET == 0
        self.capacity, self.dim, self.device = int(capacity), int(dim), device
        self.state_words = int(with_state)
        if alloc is None:
            def alloc(shape, dtype):
                return torch.empty(shape, dtype=dtype, device=device)
        self.keys = alloc((capacity,), torch.int32)
        self.keys.fill_(-1)  # 0xFFFFFFFF = empty
        self.rows = alloc((capacity, dim), torch.float32)
        self.state = None
        if self.state_words:
            self.state = alloc((capacity, self.state_words * dim), torch.float32)      # [cap][words][dim]
            self.state.zero_()
        # optional bf16 mirror of the rows (read by bf16-only pulls: half the row bytes)
        self.shadow = alloc((capacity, dim), torch.bfloat16) if with_shadow else None

    def ptrs(self):
        return (self.keys.data_ptr(), self.rows.data_ptr(), 0 if self.state is None else self.state.data_ptr(),
                0 if self.shadow is None else self.shadow.data_ptr())

    def nbytes(self):
        n = self.keys.numel() * 4 + self.rows.numel() * 4
        n += self.shadow.numel() * 2 if self.shadow is not None else 0
        return n + (self.state.numel() * 4 if self.state is not None else 0)


class ShardedTable:
    """All shards of the table as seen from one worker process.

    ``shard_ptrs``: list of (keys_ptr, rows_ptr, state_ptr) for every shard, in shard order.
    Use :meth:`local` for single-process tables (1..16 shards on one GPU)."""

    def __init__(self, hashfrag: HashFrag, capacity: int, dim: int, shard_ptrs, device, seed=0x5EED,
                 init_scale=None, owned=None, rank=0, state_words=None, init_kind=0):
        if hashfrag.num_nodes != len(shard_ptrs):
            raise ValueError("hashfrag has %d nodes but %d shards given" % (hashfrag.num_nodes, len(shard_ptrs)))
        if len(shard_ptrs) > _lib.MAX_SHARDS:
            raise ValueError("at most %d shards" % _lib.MAX_SHARDS)
        self.hashfrag, self.capacity, self.dim = hashfrag, int(capacity), int(dim)
        self.device, self.rank = device, rank
        self.num_shards = len(shard_ptrs)
        self.owned = owned or []          # DeviceShard objects this process owns (keeps them alive)
        self.frag_map_dev = hashfrag.device_map(device)
        self.init_scale = float(init_scale if init_scale is not None else 1.0 / dim)
        self.seed = int(seed)
        v = _lib.SsbTableView()
        for i, ptrs in enumerate(shard_ptrs):
            k, r, s = ptrs[0], ptrs[1], ptrs[2]
            sh = ptrs[3] if len(ptrs) > 3 else 0
            v.keys[i], v.rows[i], v.state[i], v.shadow[i] = k, r, (s or None), (sh or None)
        v.frag_map = self.frag_map_dev.data_ptr()
        v.frag_num = hashfrag.frag_num
        v.capacity = self.capacity
        v.dim = self.dim
        v.num_shards = self.num_shards
        v.seed = self.seed
        v.init_scale = self.init_scale
        self.has_state = all(p[2] for p in shard_ptrs)
        # optimizer-state words per weight ([cap][words][dim]); peers must have allocated the same
        self.state_words = int(state_words if state_words is not None else
                               (self.owned[0].state_words if self.owned else (1 if self.has_state else 0)))
        self.init_kind = int(init_kind)             # ops.rules.INIT_*: how a pull creates a missing row
        v.state_words = max(self.state_words, 1)
        v.init_kind = self.init_kind
        self.view = v
        self.has_shadow = all(len(p) > 3 and p[3] for p in shard_ptrs)
        self.stats = torch.zeros(4, dtype=torch.int32, device=device)  # [inserts, misses/full, -, -]
        # (base ptr, total 256-byte rows, rows per shard segment, first mirror row in a segment) when the
        # bf16 mirrors of all shards lie in ONE virtual range (used by the fused-pull tensor-core kernel)
        self.mirror_map = None
        # worker-side key -> slot cache (enable_slot_cache): None = every pull probes the owner's key bucket
        self.slot_cache, self.slot_cache_bits, self.slot_cache_dense = None, 0, False
        if self.has_shadow and self.num_shards == 1 and self.dim == 128:
            self.mirror_map = (shard_ptrs[0][3], self.capacity, 0, 0)

    # -- construction --------------------------------------------------------------------------
    @classmethod
    def local(cls, capacity, dim, device, num_shards=1, frag_num=None, with_state=False, with_shadow=False, **kw):
        hf = HashFrag(num_shards, frag_num or max(num_shards, 1024 if num_shards > 1 else 1))
        shards = [DeviceShard(capacity, dim, device, with_state, with_shadow=with_shadow) for _ in range(num_shards)]
        return cls(hf, capacity, dim, [s.ptrs() for s in shards], device, owned=shards, **kw)

    def _view_ref(self):
        return C.byref(self.view)

    # -- pull ----------------------------------------------------------------------------------
    def pull(self, keys, n_dev=None, rows_out=None, gslot_out=None, grads_zero=None, insert=True, n=None,
             rows_bf16=None, want_fp32=True):
        """Route + probe/insert + gather in one kernel. ``keys`` int32 (uint32 bits) [n]."""
        check_cuda(keys)
        n = keys.numel() if n is None else n
        if gslot_out is None:
            gslot_out = torch.empty(n, dtype=torch.int32, device=keys.device)
        if rows_out is None and want_fp32:
            rows_out = torch.empty((n, self.dim), dtype=torch.float32, device=keys.device)
        _lib.call("ssb_pull_cached", self._view_ref(), keys.data_ptr(), n, _lib.ptr(n_dev), gslot_out.data_ptr(),
                  _lib.ptr(rows_out), _lib.ptr(rows_bf16), _lib.ptr(grads_zero), 1 if insert else 0,
                  self.stats.data_ptr(), _lib.ptr(self.slot_cache), self._cache_bits_arg(), _lib.stream_ptr())
        return rows_out, gslot_out

    def _cache_bits_arg(self):
        return -self.slot_cache_bits if self.slot_cache_dense else self.slot_cache_bits

    def enable_slot_cache(self, expected_keys, max_bits=26, dense=False):
        """Give this worker a direct-mapped key -> (shard, slot) cache (8-byte entries, own device memory).
        Keys it has pulled before then skip the probe of the owner's key bucket — for a remote owner one
        of the two dependent NVLink round trips of a pull. ``dense=True``: the keys are small integers
        ``< expected_keys`` (word ids), the entry index is the key itself and no two keys share an entry;
        otherwise >= 2x ``expected_keys`` entries indexed by a multiplicative hash (keys that collide
        evict each other). Slots never move and are never freed (table.cu header), so entries cannot go stale."""
        want = max(int(expected_keys), 1) if dense else 2 * max(int(expected_keys), 1)
        bits = max(10, min(int(max_bits), (want - 1).bit_length()))
        if self.slot_cache is None or self.slot_cache_bits != bits or self.slot_cache_dense != bool(dense):
            self.slot_cache = torch.full((1 << bits,), -1, dtype=torch.int64, device=self.device)
            self.slot_cache_bits, self.slot_cache_dense = bits, bool(dense)
        return self.slot_cache

    def resolve(self, keys, insert=True):
        """Keys -> global slots without gathering the rows."""
        n = keys.numel()
        gslot = torch.empty(n, dtype=torch.int32, device=keys.device)
        _lib.call("ssb_pull_cached", self._view_ref(), keys.data_ptr(), n, None, gslot.data_ptr(), None, None, None,
                  1 if insert else 0, self.stats.data_ptr(), _lib.ptr(self.slot_cache), self._cache_bits_arg(),
                  _lib.stream_ptr())
        return gslot

    def gather_slots(self, gslot, rows_out=None, n_dev=None, n=None):
        n = gslot.numel() if n is None else n
        if rows_out is None:
            rows_out = torch.empty((n, self.dim), dtype=torch.float32, device=gslot.device)
        _lib.call("ssb_gather_slots", self._view_ref(), gslot.data_ptr(), n, _lib.ptr(n_dev),
                  rows_out.data_ptr(), _lib.stream_ptr())
        return rows_out

    # -- push ----------------------------------------------------------------------------------
    def push_sgd(self, gslot, grads, lr, n_dev=None, zero_grads=True, n=None):
        n = gslot.numel() if n is None else n
        _lib.call("ssb_push_sgd", self._view_ref(), gslot.data_ptr(), grads.data_ptr(), n, _lib.ptr(n_dev),
                  float(lr), 1 if zero_grads else 0, _lib.stream_ptr())

    def push_adagrad(self, gslot, grads, lr, eps=1e-8, n_dev=None, zero_grads=True, n=None):
        if not self.has_state:
            raise ValueError("table was created without AdaGrad state")
        n = gslot.numel() if n is None else n
        _lib.call("ssb_push_adagrad", self._view_ref(), gslot.data_ptr(), grads.data_ptr(), n, _lib.ptr(n_dev),
                  float(lr), float(eps), 1 if zero_grads else 0, _lib.stream_ptr())

    def push_rule(self, rule, gslot, grads, n_dev=None, zero_grads=True, n=None, only_shard=-1, **hyper):
        """Push with a named device-side update rule (ops.rules / include/ssb200_rule.cuh): ``rule`` is a name
        ("sgd", "adagrad", "ftrl", or one loaded with ``Rule.from_source``) or a ``Rule``; ``hyper`` = lr, eps, a, b,
        c, step. ref: PushAccessMethod::apply_push_value, sparse_access_method.h:36-46."""
        from . import rules

        r = rule if isinstance(rule, rules.Rule) else rules.Rule.get(rule)
        if r.state_words > self.state_words or (r.state_words and not self.has_state):
            raise ValueError("rule %r needs with_state=%d, the table has %d" % (r.name, r.state_words, self.state_words))
        r.push(self, gslot, grads, r.hyper(**hyper), n_dev=n_dev, zero_grads=zero_grads, n=n, only_shard=only_shard)

    # -- dump / load ---------------------------------------------------------------------------
    def dump_shard(self, shard: DeviceShard, with_rows=True):
        """Compact the live entries of an owned shard -> (keys uint32 np, rows float32 np)."""
        cap = shard.capacity
        n_live = torch.zeros(1, dtype=torch.int32, device=shard.device)
        # first pass: count (keys only) so the row buffer is sized exactly
        out_keys = torch.empty(cap, dtype=torch.int32, device=shard.device)
        _lib.call("ssb_dump_compact", shard.keys.data_ptr(), shard.rows.data_ptr(), cap, shard.dim,
                  out_keys.data_ptr(), None, n_live.data_ptr(), cap, _lib.stream_ptr())
        n = int(n_live.item())
        if not with_rows:
            return out_keys[:n].cpu().numpy().view(np.uint32), None
        out_rows = torch.empty((max(n, 1), shard.dim), dtype=torch.float32, device=shard.device)
        _lib.call("ssb_dump_compact", shard.keys.data_ptr(), shard.rows.data_ptr(), cap, shard.dim,
                  out_keys.data_ptr(), out_rows.data_ptr(), n_live.data_ptr(), n, _lib.stream_ptr())
        torch.cuda.synchronize(shard.device)
        return out_keys[:n].cpu().numpy().view(np.uint32), out_rows[:n].cpu().numpy()

    def iter_shard_chunks(self, shard: DeviceShard, with_state=False, max_bytes=256 << 20):
        """Stream the live entries of an owned shard in bounded pieces: yields (keys uint32 np, rows float32 np,
        state float32 np | None). One fused compaction kernel per piece copies a slot's row and optimizer-state
        row to the same index; device and host staging are sized by ``max_bytes``, not by the shard (config 4
        is 6.4 GB of rows + 6.4 GB of state per shard)."""
        want_state = bool(with_state) and shard.state is not None
        sdim = shard.dim * max(shard.state_words, 1)           # floats of optimizer state per entry
        per_slot = 4 + shard.dim * 4 + (sdim * 4 if want_state else 0)
        chunk = max(_lib.BUCKET, (int(max_bytes) // per_slot) // _lib.BUCKET * _lib.BUCKET)
        chunk = min(chunk, shard.capacity)
        dev = shard.device
        n_live = torch.zeros(1, dtype=torch.int32, device=dev)
        out_keys = torch.empty(chunk, dtype=torch.int32, device=dev)
        out_rows = torch.empty((chunk, shard.dim), dtype=torch.float32, device=dev)
        out_state = torch.empty((chunk, sdim), dtype=torch.float32, device=dev) if want_state else None
        for lo in range(0, shard.capacity, chunk):
            hi = min(lo + chunk, shard.capacity)
            _lib.call("ssb_dump_compact_range", shard.keys.data_ptr(), shard.rows.data_ptr(),
                      shard.state.data_ptr() if want_state else None, shard.capacity, shard.dim, lo, hi,
                      out_keys.data_ptr(), out_rows.data_ptr(), _lib.ptr(out_state), n_live.data_ptr(), chunk, sdim,
                      _lib.stream_ptr())
            n = int(n_live.item())
            if n == 0:
                continue
            yield (out_keys[:n].cpu().numpy().view(np.uint32), out_rows[:n].cpu().numpy(),
                   out_state[:n].cpu().numpy() if want_state else None)

    def load(self, keys, rows, state=None):
        """Insert (key,row[,state]) entries (checkpoint resume; not in the reference). ``state`` is
        [n, state_words * dim]."""
        check_cuda(keys, rows)
        if state is not None and state.numel() != keys.numel() * self.dim * max(self.state_words, 1):
            raise ValueError("state has %d floats per entry, the table keeps %d"
                             % (state.numel() // max(keys.numel(), 1), self.dim * max(self.state_words, 1)))
        _lib.call("ssb_table_load", self._view_ref(), keys.data_ptr(), rows.data_ptr(), _lib.ptr(state),
                  keys.numel(), self.stats.data_ptr(), _lib.stream_ptr())

    def size(self):
        """Number of live keys across owned shards."""
        return int(sum(((s.keys != -1) & (s.keys != -2)).sum().item() for s in self.owned))
