"""Build the sharded table for a job: one server shard per GPU, every shard visible to every
worker.  Single process => plain local shard; ``torch.distributed`` initialised (one process
per GPU, NCCL) => VMM peer mapping over NVLink (``peer.PeerMemory``).

Node numbering follows the reference (master 0, servers 1..S:
/root/reference/src/core/system/ServerWorkerRoute.h:17-31); shard index = node id - 1 = rank.
"""
from __future__ import annotations

import math

import torch

from .. import _lib
from ..ops.table import DeviceShard, ShardedTable, round_capacity
from .hashfrag import HashFrag
from .peer import PeerMemory


def shard_layout(capacity, dim, with_state, with_shadow=False):
    """Byte offsets of keys / rows / state inside a shard segment.  rows_off and the total are
    multiples of the row pitch so the mapped range is one 2-D [rows, dim] tensor for TMA."""
    pitch = dim * 4
    unit = math.lcm(4096, pitch)
    keys_bytes = capacity * 4
    rows_off = (keys_bytes + unit - 1) // unit * unit
    rows_bytes = capacity * pitch
    state_off = rows_off + rows_bytes
    shadow_off = state_off + (rows_bytes if with_state else 0)
    total = shadow_off + (rows_bytes // 2 if with_shadow else 0)
    return {"keys_off": 0, "rows_off": rows_off, "state_off": state_off if with_state else None,
            "shadow_off": shadow_off if with_shadow else None, "total": total, "pitch": pitch}


def make_table(rows_per_shard, dim, device, with_state=False, init_scale=None, frag_num=1024, seed=0x5EED,
               load_factor=0.5, force_peer=False, with_shadow=False):
    import torch.distributed as dist

    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    capacity = round_capacity(rows_per_shard, load_factor)
    hf = HashFrag(world, max(frag_num, world))
    if world == 1 and not force_peer:
        shard = DeviceShard(capacity, dim, device, with_state, with_shadow=with_shadow)
        return ShardedTable(hf, capacity, dim, [shard.ptrs()], device, seed=seed, init_scale=init_scale,
                            owned=[shard], rank=0)
    lay = shard_layout(capacity, dim, with_state, with_shadow)
    pm = PeerMemory(lay["total"], device, align_to=lay["pitch"])
    seg = pm.local_tensor()

    def alloc_at(off, shape, dtype):
        n = 1
        for s in shape:
            n *= s
        esz = 2 if dtype == torch.bfloat16 else 4
        return seg[off:off + n * esz].view(dtype).view(*shape)

    class _Shard(DeviceShard):
        pass

    offs = {"keys": lay["keys_off"], "rows": lay["rows_off"], "state": lay["state_off"], "shadow": lay["shadow_off"]}
    order = iter(["keys", "rows"] + (["state"] if with_state else []) + (["shadow"] if with_shadow else []))
    shard = DeviceShard(capacity, dim, device, with_state, with_shadow=with_shadow,
                        alloc=lambda shape, dtype: alloc_at(offs[next(order)], shape, dtype))
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()  # every shard is initialised before anyone probes it
    ptrs = []
    for r in range(world):
        b = pm.segment_ptr(r)
        ptrs.append((b + lay["keys_off"], b + lay["rows_off"], (b + lay["state_off"]) if with_state else 0,
                     (b + lay["shadow_off"]) if with_shadow else 0))
    table = ShardedTable(hf, capacity, dim, ptrs, device, seed=seed, init_scale=init_scale, owned=[shard],
                         rank=rank)
    table.peer_memory = pm
    table.layout = lay
    if with_shadow and dim == 128 and pm.stride % 256 == 0 and lay["shadow_off"] % 256 == 0:
        table.mirror_map = (pm.base, world * pm.stride // 256, pm.stride // 256, lay["shadow_off"] // 256)
    return table


class Inbox:
    """Owner-side gradient inboxes for the one-way multi-GPU push (csrc/kernels/table.cu,
    push_send_kernel): every rank owns two inboxes (step parity) in peer-mapped memory, each split
    into one private sub-region per sender.  A push ships (slot, grad row) into the sender's own
    sub-region on each owner — space is claimed with local atomics, no remote atomic exists on this
    path — then publishes its row counts and raises hand-off flags in peer memory; an owner waits
    for every sender's flag (no collective on the hot path) and applies all sub-regions to its own
    shard in one launch at local-HBM speed.  This is the reference's "send (key, grad)* to the
    server, the server applies" (global_push_access.h:123-149, system/server/init.h:116-132) carried
    by NVLink stores instead of ZeroMQ frames.  ``bf16_rows`` ships the gradient rows as bf16 (half
    the NVLink bytes; the update stays fp32 on fp32 rows and accumulators).  ``barrier`` selects the
    hand-off: "flags" (default) or "nccl" (a 4-byte all-reduce).  ``capacity`` = rows per owner
    inbox; a sender may use ``capacity // world`` of them (excess rows are counted in ``overflow``
    and dropped)."""

    def __init__(self, table: ShardedTable, capacity: int, device, bf16_rows: bool = False, barrier: str = "flags"):
        import ctypes as C

        import torch.distributed as dist

        self.table, self.device = table, device
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.seg_rows = max(64, (int(capacity) + self.world - 1) // self.world)     # rows per (owner, sender)
        self.capacity = self.seg_rows * self.world
        self.bf16_rows, self.barrier = bool(bf16_rows), barrier
        D = table.dim
        esz = 2 if self.bf16_rows else 4
        rows_b = self.capacity * D * esz
        slots_off = rows_b
        count_off = slots_off + self.capacity * 4
        per = (count_off + 256 + 4095) // 4096 * 4096
        flags_off = 2 * per
        self.pm = PeerMemory(flags_off + 4096, device)
        seg = self.pm.local_tensor()
        seg.zero_()
        self.views, self.rows_l, self.slots_l, self.count_l = [], [], [], []
        for par in range(2):
            # the view a SENDER uses: its private sub-region on every owner
            v = _lib.SsbInboxView()
            for r in range(self.world):
                b = self.pm.segment_ptr(r) + par * per
                v.rows[r] = b + self.rank * self.seg_rows * D * esz
                v.slots[r] = b + slots_off + self.rank * self.seg_rows * 4
                v.count[r] = b + count_off + self.rank * 4
                v.flags[r] = self.pm.segment_ptr(r) + flags_off
            v.capacity = self.seg_rows
            self.views.append(v)
            # what the OWNER applies: all sub-regions of its own inbox
            base = par * per
            rows = seg[base:base + rows_b]
            self.rows_l.append(rows.view(torch.bfloat16 if self.bf16_rows else torch.float32).view(self.capacity, D))
            self.slots_l.append(seg[base + slots_off:base + slots_off + self.capacity * 4].view(torch.int32))
            self.count_l.append(seg[base + count_off:base + count_off + 4 * self.world].view(torch.int32))
        self.flags_l = seg[flags_off:flags_off + 4 * self.world].view(torch.int32)
        shard = table.owned[0]
        self.local = ShardedTable(HashFrag(1, 1), table.capacity, D, [shard.ptrs()], device, seed=table.seed,
                                  init_scale=table.init_scale, owned=[shard])
        self.overflow = torch.zeros(1, dtype=torch.int32, device=device)
        self.cursor = torch.zeros(16, dtype=torch.int32, device=device)     # rows claimed per owner this step
        self.token = torch.zeros(1, dtype=torch.int32, device=device)
        self.parity = 0
        self.step = 0
        self._C = C
        torch.cuda.synchronize(device)
        dist.barrier()

    def push(self, gslot, grads, lr, eps=1e-8, adagrad=True, n_dev=None, n=None):
        import torch.distributed as dist

        C = self._C
        n = gslot.numel() if n is None else n
        par = self.parity
        self.parity ^= 1
        self.step += 1
        sp = _lib.stream_ptr()
        view = C.byref(self.views[par])
        _lib.call("ssb_push_send2", self.table._view_ref(), view, gslot.data_ptr(), grads.data_ptr(), n,
                  _lib.ptr(n_dev), self.rank, float(lr), float(eps), 1 if adagrad else 0, 1 if self.bf16_rows else 0,
                  self.overflow.data_ptr(), self.cursor.data_ptr(), sp)
        if self.barrier == "flags":
            _lib.call("ssb_inbox_publish", view, self.world, self.rank, self.cursor.data_ptr(), self.step, sp)
            _lib.call("ssb_flags_wait", self.flags_l.data_ptr(), self.world, self.step, sp)
        else:
            _lib.call("ssb_inbox_publish", view, self.world, self.rank, self.cursor.data_ptr(), 0, sp)
            dist.all_reduce(self.token)      # every rank's shipment has landed in my inbox
        if self.bf16_rows:
            _lib.call("ssb_push_bf16_seg", self.local._view_ref(), self.slots_l[par].data_ptr(), self.rows_l[par].data_ptr(),
                      self.seg_rows, self.world, self.count_l[par].data_ptr(), float(lr), float(eps), 1 if adagrad else 0, sp)
        else:
            S = self.seg_rows
            for r in range(self.world):
                if r == self.rank:
                    continue                 # own rows were applied directly by the send kernel
                sl, rw, cn = self.slots_l[par][r * S:(r + 1) * S], self.rows_l[par][r * S:(r + 1) * S], self.count_l[par][r:r + 1]
                if adagrad:
                    self.local.push_adagrad(sl, rw, lr, eps, n_dev=cn, zero_grads=False, n=S)
                else:
                    self.local.push_sgd(sl, rw, lr, n_dev=cn, zero_grads=False, n=S)
        # no clearing: every sender overwrites its count on its next publish (also with 0 rows)


def attach_inbox(table: ShardedTable, capacity: int, bf16_rows: bool = False, barrier: str = "flags"):
    """Give a peer-mapped table owner-side inboxes (no-op for a single process)."""
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size() == 1:
        table.inbox = None
        return None
    table.inbox = Inbox(table, capacity, table.device, bf16_rows=bf16_rows, barrier=barrier)
    return table.inbox
