"""Hash-fragment key sharding (key -> fragment -> server node).

ref: /root/reference/src/core/parameter/hashfrag.h:12-116 (BasicHashFrag) and
utils/HashFunction.h:16-24 (fmix64).  The assignment rule is reproduced exactly so fragment
maps and dumps are comparable with the reference: fragment ``i`` goes to node
``i // (frag_num // num_nodes) + 1`` clamped to ``[1, num_nodes]`` (remainder fragments land on
the last server).  Unlike the reference (hashfrag.h:36, divide-by-zero when
``frag_num < num_nodes``) the constructor rejects that configuration.

Wire form (hashfrag.h:55-85): ``int32 num_nodes, int32 num_frags, uint32 map[num_frags]``.
"""
from __future__ import annotations

import struct

import numpy as np

_M1 = np.uint64(0xFF51AFD7ED558CCD)
_M2 = np.uint64(0xC4CEB9FE1A85EC53)
_S33 = np.uint64(33)


def fmix64(x):
    """MurmurHash3 64-bit finaliser, vectorised over a numpy array (or scalar)."""
    x = np.asarray(x).astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        x ^= x >> _S33
        x *= _M1
        x ^= x >> _S33
        x *= _M2
        x ^= x >> _S33
    return x


class HashFrag:
    def __init__(self, num_nodes: int, frag_num: int, frag_map=None):
        if num_nodes < 1:
            raise ValueError("num_nodes must be >= 1")
        if frag_num < num_nodes:
            raise ValueError("frag_num (%d) must be >= number of servers (%d)" % (frag_num, num_nodes))
        self.num_nodes = int(num_nodes)
        self.frag_num = int(frag_num)
        if frag_map is None:
            per = self.frag_num // self.num_nodes
            m = np.arange(self.frag_num, dtype=np.int64) // per + 1
            m = np.clip(m, 1, self.num_nodes)
            frag_map = m.astype(np.uint32)
        self.map = np.ascontiguousarray(frag_map, dtype=np.uint32)
        assert self.map.shape == (self.frag_num,)

    # -- routing ---------------------------------------------------------------------------
    def to_frag(self, keys):
        return (fmix64(keys) % np.uint64(self.frag_num)).astype(np.int64)

    def to_node_id(self, keys):
        """Node id (1..num_nodes) per key. ref: hashfrag.h:48-53."""
        return self.map[self.to_frag(keys)].astype(np.int32)

    def to_shard(self, keys):
        """Zero-based shard / GPU index per key."""
        return self.to_node_id(keys) - 1

    # -- wire form -------------------------------------------------------------------------
    def serialize(self) -> bytes:
        return struct.pack("<ii", self.num_nodes, self.frag_num) + self.map.tobytes()

    @classmethod
    def deserialize(cls, buf: bytes) -> "HashFrag":
        n, f = struct.unpack_from("<ii", buf, 0)
        m = np.frombuffer(buf, dtype=np.uint32, count=f, offset=8).copy()
        return cls(n, f, m)

    def device_map(self, device):
        import torch

        return torch.from_numpy(self.map.view(np.int32).copy()).to(device)

    def __repr__(self):
        return "HashFrag(num_nodes=%d, frag_num=%d)" % (self.num_nodes, self.frag_num)
