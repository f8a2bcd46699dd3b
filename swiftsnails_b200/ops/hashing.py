"""K1 — key hashing / routing / bucketing on the GPU.

ref: /root/reference/src/utils/HashFunction.h:16-24, core/parameter/hashfrag.h:48-53,
core/parameter/global_pull_access.h:58-72 (arrange_local_vals).
"""
from __future__ import annotations

import torch

from .. import _lib
from ._util import check_cuda


def fmix64(x: torch.Tensor) -> torch.Tensor:
    check_cuda(x)
    assert x.dtype == torch.int64
    out = torch.empty_like(x)
    _lib.call("ssb_fmix64", x.data_ptr(), out.data_ptr(), x.numel(), _lib.stream_ptr())
    return out


def hash_route(keys: torch.Tensor, frag_map: torch.Tensor, num_owners: int):
    """Return (owner[int32, n], counts[int32, num_owners]); owner = node_id - 1."""
    check_cuda(keys, frag_map)
    n = keys.numel()
    owner = torch.empty(n, dtype=torch.int32, device=keys.device)
    counts = torch.zeros(num_owners, dtype=torch.int32, device=keys.device)
    _lib.call("ssb_hash_route", keys.data_ptr(), n, frag_map.data_ptr(), frag_map.numel(), num_owners,
              owner.data_ptr(), counts.data_ptr(), _lib.stream_ptr())
    return owner, counts


def bucket_by_owner(keys: torch.Tensor, frag_map: torch.Tensor, num_owners: int):
    """Partition keys by owner. Returns (bucketed_keys, perm, counts, offsets) — bucketed_keys
    [offsets[o]:offsets[o+1]] are the keys owned by shard o, perm the original positions."""
    owner, counts = hash_route(keys, frag_map, num_owners)
    offsets = torch.zeros(num_owners + 1, dtype=torch.int32, device=keys.device)
    offsets[1:] = torch.cumsum(counts, 0)
    cursor = torch.zeros(num_owners, dtype=torch.int32, device=keys.device)
    bucketed = torch.empty_like(keys)
    perm = torch.empty_like(keys)
    _lib.call("ssb_bucket_scatter", keys.data_ptr(), owner.data_ptr(), keys.numel(), offsets.data_ptr(),
              cursor.data_ptr(), bucketed.data_ptr(), perm.data_ptr(), _lib.stream_ptr())
    return bucketed, perm, counts, offsets
