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
from .peer import PeerMemory, tensor_from_ptr


def shard_layout(capacity, dim, with_state):
    """Byte offsets of keys / rows / state inside a shard segment.  rows_off and the total are
    multiples of the row pitch so the mapped range is one 2-D [rows, dim] tensor for TMA."""
    pitch = dim * 4
    unit = math.lcm(4096, pitch)
    keys_bytes = capacity * 4
    rows_off = (keys_bytes + unit - 1) // unit * unit
    rows_bytes = capacity * pitch
    state_off = rows_off + rows_bytes
    total = state_off + (rows_bytes if with_state else 0)
    return {"keys_off": 0, "rows_off": rows_off, "state_off": state_off if with_state else None,
            "total": total, "pitch": pitch}


def make_table(rows_per_shard, dim, device, with_state=False, init_scale=None, frag_num=1024, seed=0x5EED,
               load_factor=0.5, force_peer=False):
    import torch.distributed as dist

    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    capacity = round_capacity(rows_per_shard, load_factor)
    hf = HashFrag(world, max(frag_num, world))
    if world == 1 and not force_peer:
        shard = DeviceShard(capacity, dim, device, with_state)
        return ShardedTable(hf, capacity, dim, [shard.ptrs()], device, seed=seed, init_scale=init_scale,
                            owned=[shard], rank=0)
    lay = shard_layout(capacity, dim, with_state)
    pm = PeerMemory(lay["total"], device, align_to=lay["pitch"])
    seg = pm.local_tensor()

    def alloc_at(off, shape, dtype):
        n = 1
        for s in shape:
            n *= s
        return seg[off:off + n * 4].view(dtype).view(*shape)

    class _Shard(DeviceShard):
        pass

    offs = {"keys": lay["keys_off"], "rows": lay["rows_off"], "state": lay["state_off"]}
    order = iter(["keys", "rows", "state"])
    shard = DeviceShard(capacity, dim, device, with_state,
                        alloc=lambda shape, dtype: alloc_at(offs[next(order)], shape, dtype))
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()  # every shard is initialised before anyone probes it
    ptrs = []
    for r in range(world):
        b = pm.segment_ptr(r)
        ptrs.append((b + lay["keys_off"], b + lay["rows_off"], (b + lay["state_off"]) if with_state else 0))
    table = ShardedTable(hf, capacity, dim, ptrs, device, seed=seed, init_scale=init_scale, owned=[shard],
                         rank=rank)
    table.peer_memory = pm
    table.layout = lay
    return table
