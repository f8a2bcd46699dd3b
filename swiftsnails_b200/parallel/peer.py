"""Peer-mapped shard memory over NVLink 5 / NVSwitch (the data-plane "transport").

ref: /root/reference/src/core/transfer/Route.h:20-112 (id -> address -> PUSH socket) and
core/system/master/init.h:74-99 (the master broadcasts the route table).  The B200 analogue of
"connect a socket to every peer" is "map every peer's shard into my address space": each rank
allocates its shard with the CUDA VMM API (``csrc/kernels/vmm.cu``), exports it as a POSIX file
descriptor, the descriptors are exchanged over Unix sockets (SCM_RIGHTS), and every rank maps
all shards side by side in one reserved virtual range.  After that, pull and push kernels
address any shard with ordinary pointers.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import socket
import struct
import threading
import uuid

import torch

from .. import _lib


class _RawCuda:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1",
                                         "data": (int(ptr), False), "version": 3}


def tensor_from_ptr(ptr, nbytes, device):
    """Zero-copy uint8 tensor over raw device memory."""
    return torch.as_tensor(_RawCuda(ptr, nbytes), device=device)


def _vmm(name, *args):
    lib = _lib.kernels()
    fn = getattr(lib, name)
    fn.restype = C.c_int
    err = fn(*args)
    if err != 0:
        kind = "CUresult %d" % (err - 100000) if err >= 100000 else "cudaError %d" % err
        raise _lib.NativeError("%s failed: %s" % (name, kind))


def vmm_granularity(device_index):
    g = C.c_uint64()
    _vmm("ssb_vmm_granularity", C.c_int(device_index), C.byref(g))
    return int(g.value)


def exchange_fds(my_fd, rank, world, tag):
    """All-to-all exchange of one file descriptor per rank over abstract Unix sockets.
    Returns {peer_rank: fd}.  Deadlock-free: a thread accepts while the caller connects."""
    name = lambda r: "\0ssb200-%s-%d" % (tag, r)
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(name(rank))
    srv.listen(world)
    got = {}

    def acceptor():
        for _ in range(world - 1):
            conn, _ = srv.accept()
            msg, fds, _, _ = socket.recv_fds(conn, 4, 1)
            got[struct.unpack("<i", msg)[0]] = fds[0]
            conn.close()

    th = threading.Thread(target=acceptor, daemon=True)
    th.start()
    from . import group as _group

    _group.current().barrier()  # every rank is listening
    for p in range(world):
        if p == rank:
            continue
        c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        c.connect(name(p))
        socket.send_fds(c, [struct.pack("<i", rank)], [my_fd])
        c.close()
    th.join(timeout=120)
    if len(got) != world - 1:
        raise RuntimeError("fd exchange incomplete: got %d of %d" % (len(got), world - 1))
    srv.close()
    return got


class PeerMemory:
    """``world`` equally-sized segments, one per rank, mapped contiguously in this process.

    ``base``            start of the reserved range (same layout, different VA, on every rank)
    ``stride``          bytes per segment (multiple of the allocation granularity)
    ``segment_ptr(r)``  device pointer of rank r's segment
    """

    def __init__(self, nbytes, device, align_to=1):
        from . import group as _group

        grp = _group.current()       # the master-run job (ZeroMQ), torch.distributed if that is what runs, or one rank
        self.device = device
        self.rank, self.world = grp.rank, grp.world
        dev_index = device.index if device.index is not None else torch.cuda.current_device()
        gran = vmm_granularity(dev_index)
        unit = math.lcm(gran, int(align_to))
        self.stride = (int(nbytes) + unit - 1) // unit * unit
        handle, fd = C.c_uint64(), C.c_int()
        _vmm("ssb_vmm_create", C.c_int(dev_index), C.c_uint64(self.stride), C.byref(handle), C.byref(fd))
        va = C.c_uint64()
        _vmm("ssb_vmm_reserve", C.c_uint64(self.stride * self.world), C.c_uint64(unit if unit % gran == 0 else gran),
             C.byref(va))
        self.base = int(va.value)
        self._handles = {self.rank: int(handle.value)}
        _vmm("ssb_vmm_map", C.c_uint64(self.base + self.rank * self.stride), C.c_uint64(self.stride),
             C.c_uint64(handle.value), C.c_int(dev_index))
        if self.world > 1:
            tag = grp.broadcast_object(uuid.uuid4().hex if self.rank == 0 else None, src=0)
            fds = exchange_fds(fd.value, self.rank, self.world, tag)
            for p, pfd in sorted(fds.items()):
                h = C.c_uint64()
                _vmm("ssb_vmm_import", C.c_int(pfd), C.byref(h))
                os.close(pfd)
                self._handles[p] = int(h.value)
                _vmm("ssb_vmm_map", C.c_uint64(self.base + p * self.stride), C.c_uint64(self.stride),
                     C.c_uint64(h.value), C.c_int(dev_index))
            grp.barrier()
        os.close(fd.value)

    def segment_ptr(self, r):
        return self.base + r * self.stride

    def local_tensor(self):
        return tensor_from_ptr(self.segment_ptr(self.rank), self.stride, self.device)

    def segment_tensor(self, r):
        return tensor_from_ptr(self.segment_ptr(r), self.stride, self.device)

    def close(self):
        for p, h in self._handles.items():
            try:
                _vmm("ssb_vmm_unmap", C.c_uint64(self.base + p * self.stride), C.c_uint64(self.stride))
                _vmm("ssb_vmm_release", C.c_uint64(h))
            except Exception:
                pass
        self._handles = {}
