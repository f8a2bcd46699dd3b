"""NCCL all-to-all pull / push — the GPU *baseline* the P2P kernels have to beat (BASELINE.json:
"an NCCL all-to-all pull/push implementation as the GPU baseline. A path that only calls NCCL is
the baseline, not the product").

It is the straightforward translation of the reference's RPC
(/root/reference/src/core/parameter/global_pull_access.h:40-120, global_push_access.h:36-149)
onto collectives: bucket keys per owner → exchange counts (host sync for the split sizes) →
``all_to_all_single`` of keys → owner-side lookup on its own shard → ``all_to_all_single`` of the
rows back → un-permute.  Push mirrors it with (key, grad) going out and the owner applying the
update.  Same table memory, same owner rule and same owner-side kernels as the P2P path (through a
one-shard local view), so the two paths are interchangeable and comparable.

Two variants: the plain one exchanges exact split sizes (one host synchronisation per pull, as ``all_to_all_single``
with uneven splits needs them on the host); ``padded=True`` gives every (sender, owner) pair a fixed bucket of
``ceil(n / world * slack)`` keys, pads with invalid keys and uses equal-split collectives — no host sync anywhere,
at the price of moving the padding; ``wire_bf16=True`` sends rows and gradients as bf16 (what the P2P path does).
Keys that do not fit their bucket are counted in ``overflow`` (and then missing from the result).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops import hashing
from ..ops.table import ShardedTable
from .hashfrag import HashFrag


class NcclShardedTable:
    """Wraps a (peer-mapped or local) ``ShardedTable`` but never touches a peer pointer."""

    def __init__(self, table: ShardedTable, padded: bool = False, slack: float = 1.25, wire_bf16: bool = False):
        self.table = table
        self.padded, self.slack, self.wire_bf16 = bool(padded), float(slack), bool(wire_bf16)
        self.overflow = torch.zeros(1, dtype=torch.int64, device=table.device)
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.dim, self.device, self.capacity = table.dim, table.device, table.capacity
        self.hashfrag = table.hashfrag
        self.has_state = table.has_state
        self.owned = table.owned
        shard = table.owned[0]
        # one-shard view of my own shard: every key routes to shard 0 = my memory
        self.local = ShardedTable(HashFrag(1, 1), table.capacity, table.dim, [shard.ptrs()], table.device,
                                  seed=table.seed, init_scale=table.init_scale, owned=[shard])
        self.stats = self.local.stats
        self._plan = None

    # ---- exchange plan for a key set ---------------------------------------------------------
    def _route(self, keys):
        bucketed, perm, counts, offsets = hashing.bucket_by_owner(keys, self.table.frag_map_dev, self.world)
        send_counts = counts.to(torch.int64)
        recv_counts = torch.empty_like(send_counts)
        if self.world > 1:
            dist.all_to_all_single(recv_counts, send_counts)
        else:
            recv_counts.copy_(send_counts)
        sc = send_counts.tolist()      # host sync: NCCL needs the split sizes on the host
        rc = recv_counts.tolist()
        recv_keys = torch.empty(sum(rc), dtype=torch.int32, device=keys.device)
        if self.world > 1:
            dist.all_to_all_single(recv_keys, bucketed, rc, sc)
        else:
            recv_keys.copy_(bucketed)
        return {"perm": perm.long(), "sc": sc, "rc": rc, "recv_keys": recv_keys}

    # ---- padded variant: fixed buckets, equal-split collectives, no host synchronisation ----------------------------
    def _route_padded(self, keys, n):
        W = self.world
        cap = (int(n / W * self.slack) + 63) // 64 * 64 + 64
        bucketed, perm, counts, offsets = hashing.bucket_by_owner(keys, self.table.frag_map_dev, W)
        pos = torch.arange(n, device=keys.device, dtype=torch.int64)
        off64 = offsets.to(torch.int64)
        live = pos < off64[W]              # invalid (padding) keys are not routed: the bucketed list is shorter than n
        owner = torch.searchsorted(off64[1:W + 1].contiguous(), pos, right=True).clamp_(max=W - 1)
        within = pos - off64[owner]
        ok = live & (within < cap)
        self.overflow += (live & ~ok).sum()
        dst = torch.where(ok, owner * cap + within, torch.full_like(pos, W * cap))      # overflow -> a scratch slot
        send = torch.full((W * cap + 1,), -1, dtype=torch.int32, device=keys.device)
        send[dst] = bucketed
        send = send[:W * cap]
        recv = torch.empty_like(send)
        if W > 1:
            dist.all_to_all_single(recv, send)
        else:
            recv.copy_(send)
        # positions past the routed keys hold no permutation entry: send them to a scratch row (index n)
        perm_safe = torch.where(live, perm.long(), torch.full_like(pos, n))
        return {"perm": perm_safe, "live": live, "dst": dst, "ok": ok, "cap": cap, "recv_keys": recv}

    def _pull_padded(self, keys, n, rows_out, rows_bf16, grads_zero, insert, want_fp32):
        plan = self._route_padded(keys[:n], n)
        W, cap = self.world, plan["cap"]
        wire = torch.bfloat16 if self.wire_bf16 else torch.float32
        if self.wire_bf16:
            rows_owner = torch.empty((W * cap, self.dim), dtype=torch.bfloat16, device=keys.device)
            self.local.pull(plan["recv_keys"], rows_bf16=rows_owner, want_fp32=False, insert=insert)
        else:
            rows_owner, _ = self.local.pull(plan["recv_keys"], insert=insert)
        back = torch.empty((W * cap + 1, self.dim), dtype=wire, device=keys.device)
        if W > 1:
            dist.all_to_all_single(back[:W * cap], rows_owner)
        else:
            back[:W * cap].copy_(rows_owner)
        back[W * cap].zero_()
        mine = back.index_select(0, plan["dst"])            # bucketed order
        if rows_out is None and want_fp32:
            rows_out = torch.empty((n, self.dim), dtype=torch.float32, device=keys.device)
        tmp = torch.zeros((n + 1, self.dim), dtype=wire, device=keys.device)          # row n = scratch
        tmp.index_copy_(0, plan["perm"], mine)
        if rows_out is not None:
            rows_out[:n].copy_(tmp[:n])
        if rows_bf16 is not None:
            rows_bf16[:n].copy_(tmp[:n])
        if grads_zero is not None:
            grads_zero[:n].zero_()
        self._plan = (plan, n)
        return rows_out

    def _push_padded(self, grads, n, apply):
        plan, n_plan = self._plan
        assert n_plan == n, "push must follow the pull of the same key set"
        W, cap = self.world, plan["cap"]
        wire = torch.bfloat16 if self.wire_bf16 else torch.float32
        g = grads[:n].index_select(0, plan["perm"].clamp(max=n - 1)).to(wire) * plan["live"][:, None].to(wire)
        send = torch.zeros((W * cap + 1, self.dim), dtype=wire, device=grads.device)
        send[plan["dst"]] = g
        recv = torch.empty((W * cap, self.dim), dtype=wire, device=grads.device)
        if W > 1:
            dist.all_to_all_single(recv, send[:W * cap])
        else:
            recv.copy_(send[:W * cap])
        gslot = self.local.resolve(plan["recv_keys"], insert=True)          # padding keys resolve to the invalid slot: skipped
        apply(gslot, recv.to(torch.float32))

    def pull(self, keys, n_dev=None, rows_out=None, gslot_out=None, grads_zero=None, insert=True, n=None,
             rows_bf16=None, want_fp32=True):
        n = keys.numel() if n is None else n
        if self.padded:
            # n_dev (the live prefix of `keys`) is honoured without reading it on the host: the tail is replaced by
            # invalid keys, which are routed like padding and come back as zero rows
            k = keys[:n]
            if n_dev is not None:
                k = torch.where(torch.arange(n, device=keys.device) < n_dev.to(torch.int64), k, torch.full_like(k, -1))
            return self._pull_padded(k, n, rows_out, rows_bf16, grads_zero, insert, want_fp32), gslot_out
        if n_dev is not None:
            n = min(n, int(n_dev.item()))
        k = keys[:n]
        plan = self._route(k)
        rows_owner, _ = self.local.pull(plan["recv_keys"], insert=insert)
        back = torch.empty((n, self.dim), dtype=torch.float32, device=keys.device)
        if self.world > 1:
            dist.all_to_all_single(back, rows_owner, plan["sc"], plan["rc"])
        else:
            back.copy_(rows_owner)
        if rows_out is None and want_fp32:
            rows_out = torch.empty((n, self.dim), dtype=torch.float32, device=keys.device)
        if rows_out is not None:
            rows_out[:n].index_copy_(0, plan["perm"], back)
        if rows_bf16 is not None:
            rows_bf16[:n].index_copy_(0, plan["perm"], back.to(torch.bfloat16))
        if grads_zero is not None:
            grads_zero[:n].zero_()
        self._plan = (plan, n)
        return rows_out, gslot_out

    def _push(self, grads, n, apply):
        if self.padded:
            return self._push_padded(grads, n, apply)
        plan, n_plan = self._plan
        assert n_plan == n, "push must follow the pull of the same key set"
        g = grads[:n].index_select(0, plan["perm"])
        recv = torch.empty((sum(plan["rc"]), self.dim), dtype=torch.float32, device=grads.device)
        if self.world > 1:
            dist.all_to_all_single(recv, g, plan["rc"], plan["sc"])
        else:
            recv.copy_(g)
        gslot = self.local.resolve(plan["recv_keys"], insert=True)
        apply(gslot, recv)

    def push_sgd(self, gslot, grads, lr, n_dev=None, zero_grads=True, n=None):
        n = self._plan[1]
        self._push(grads, n, lambda s, g: self.local.push_sgd(s, g, lr, zero_grads=False))
        if zero_grads:
            grads[:n].zero_()

    def push_adagrad(self, gslot, grads, lr, eps=1e-8, n_dev=None, zero_grads=True, n=None):
        n = self._plan[1]
        self._push(grads, n, lambda s, g: self.local.push_adagrad(s, g, lr, eps, zero_grads=False))
        if zero_grads:
            grads[:n].zero_()
