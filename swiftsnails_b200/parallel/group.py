"""The job's process group: who am I, how many are we, and the few host-side collectives the set-up needs.

ref: /root/reference/src/core/system/master/init.h:21-171 (the master waits for every node, then hands out the route
table), system/node_init.h:16-152 (node side), master/terminate.h + server/terminate.h + worker/terminate.h (shutdown).

A GPU job of this framework is run by the MASTER, as in the reference: every GPU process (server shard g + worker g)
registers with ``swift_master`` over ZeroMQ and gets its rank, the route table and the hash-fragment map from it
(:class:`MasterGroup`, native code in ``csrc/core/system/group_exchange.h``). The handful of host-side collectives the
set-up needs — exchange of the socket names that carry the VMM file descriptors, barriers around table creation, the
max over ranks of a device-timed interval — are all-gathers of small byte strings THROUGH the master. Nothing on the
training path uses them: pulls and pushes are NVLink loads / stores / atomics on peer-mapped memory and the push
hand-off is a flag in peer memory.

``torch.distributed`` is not needed for any of this; :class:`TorchGroup` wraps it for the NCCL baseline arm
(``--impl nccl``) and for callers that have already initialised it. :func:`current` returns the active group
(a one-rank group when nothing was set up)."""
from __future__ import annotations

import ctypes as C
import os
import pickle
import subprocess
import tempfile
import time

from .. import _lib

_current = None


class Group:
    """One rank's handle on the job. Subclasses implement :meth:`all_gather_bytes`; the rest derives from it."""
    rank, world, kind = 0, 1, "single"

    def all_gather_bytes(self, data: bytes) -> list:
        return [data]

    def all_gather_object(self, obj) -> list:
        return [pickle.loads(b) for b in self.all_gather_bytes(pickle.dumps(obj))]

    def broadcast_object(self, obj, src=0):
        return self.all_gather_object(obj if self.rank == src else None)[src]

    def barrier(self):
        self.all_gather_bytes(b"")

    def all_reduce(self, value, op="sum"):
        """Host-side reduction of one Python number over the ranks (sum | max | min)."""
        vals = self.all_gather_object(value)
        return {"sum": sum, "max": max, "min": min}[op](vals)

    def hashfrag(self, frag_num=None):
        """The job's fragment map (parallel.hashfrag.HashFrag)."""
        from .hashfrag import HashFrag

        return HashFrag(self.world, max(frag_num or 1024, self.world))

    def finish(self, dump=None):
        """End of the job: run ``dump`` (a callable writing this rank's shard) at the point the protocol says."""
        if dump is not None:
            dump()

    def close(self):
        pass


class TorchGroup(Group):
    """An already-initialised ``torch.distributed`` job (the NCCL baseline arm)."""
    kind = "torch"

    def __init__(self):
        import torch.distributed as dist

        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        self.rank, self.world = dist.get_rank(), dist.get_world_size()

    def all_gather_bytes(self, data: bytes) -> list:
        import torch.distributed as dist

        out = [None] * self.world
        dist.all_gather_object(out, data)
        return out

    def barrier(self):
        import torch.distributed as dist

        dist.barrier()


class MasterGroup(Group):
    """A GPU rank registered with ``swift_master`` (ZeroMQ): rank, route, fragment map and collectives come from
    the master; shutdown is WORKER_FINISH_WORK -> SERVER_TOLD_TO_TERMINATE -> dump -> ack."""
    kind = "master"

    def __init__(self, master_addr: str, listen_addr: str = "", init_timeout=120, op_timeout=600,
                 master_proc=None):
        lib = _lib.core()
        lib.ssb_node_start.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        lib.ssb_node_hashfrag.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        lib.ssb_node_exchange.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_long, C.c_void_p, C.c_int]
        lib.ssb_node_exchange.restype = C.c_long
        lib.ssb_node_finish.argtypes = [C.c_int]
        self._lib, self.op_timeout, self.master_proc = lib, int(op_timeout), master_proc
        self.master_addr = master_addr
        r = lib.ssb_node_start(master_addr.encode(), listen_addr.encode(), int(init_timeout))
        if r < 0:
            raise RuntimeError("registration with the master at %s failed (%d)" % (master_addr, r))
        self.rank, self.world = int(r), int(lib.ssb_node_world())
        self._finished = False

    def all_gather_bytes(self, data: bytes) -> list:
        cap = 1 << 20
        while True:
            out = C.create_string_buffer(cap)
            lens = (C.c_long * self.world)()
            n = self._lib.ssb_node_exchange(data, len(data), out, cap, lens, self.op_timeout)
            if n == -3:                       # cannot happen twice: the reply of a round is delivered once
                raise RuntimeError("group exchange: reply larger than %d bytes" % cap)
            if n < 0:
                raise RuntimeError("group exchange through the master failed (%d)" % n)
            parts, off = [], 0
            for r in range(self.world):
                parts.append(out.raw[off:off + lens[r]])
                off += lens[r]
            return parts

    def all_gather_object(self, obj) -> list:
        return [pickle.loads(b) for b in self.all_gather_bytes(pickle.dumps(obj))]

    def hashfrag(self, frag_num=None):
        """The map the master built from its own ``frag_num`` (NODE_ASKFOR_HASHFRAG); ``frag_num`` is ignored."""
        import numpy as np

        from .hashfrag import HashFrag

        info = (C.c_int * 2)()
        n = self._lib.ssb_node_hashfrag(info, None, 0)
        m = np.zeros(n, dtype=np.uint32)
        if self._lib.ssb_node_hashfrag(info, m.ctypes.data, n) != n:
            raise RuntimeError("fragment map not available")
        return HashFrag(int(info[0]), int(info[1]), frag_map=m)

    def finish(self, dump=None):
        if self._finished:
            return
        self._finished = True
        rc = self._lib.ssb_node_finish(self.op_timeout)
        if rc != 0:
            raise RuntimeError("the master did not complete the termination sequence (%d)" % rc)
        try:
            if dump is not None:
                dump()
        finally:
            self._lib.ssb_node_dump_done()

    def close(self):
        if not self._finished:
            self.finish()
        self._lib.ssb_node_stop()
        if self.master_proc is not None:
            try:
                self.master_proc.wait(timeout=30)
            except subprocess.TimeoutExpired:
                self.master_proc.kill()
            self.master_proc = None


def current() -> Group:
    """The active group: what :func:`init` / :func:`set_current` installed, else an initialised torch.distributed
    job, else a single rank."""
    global _current
    if _current is not None:
        return _current
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return TorchGroup()
    except ImportError:
        pass
    return Group()


def set_current(g):
    global _current
    _current = g
    return g


def spawn_master(listen_addr: str, world: int, frag_num: int = 1024, extra=None, log_path=None):
    """Start ``swift_master`` (csrc/apps/swift_master.cc) for a GPU job of ``world`` ranks. Returns the Popen."""
    from .. import build as _build

    exe = os.path.join(_build.BIN_DIR, "swift_master")
    if not os.path.exists(exe):
        _build.build_core()
    conf = tempfile.NamedTemporaryFile("w", suffix=".conf", prefix="ssb200_master_", delete=False)
    conf.write("listen_addr: %s\nexpected_node_num: %d\nfrag_num: %d\ncolocated_workers: 1\nmaster_time_out: 300\n"
               "master_longest_alive_duration: 86400\n" % (listen_addr, world, max(frag_num, world)))
    for k, v in (extra or {}).items():
        conf.write("%s: %s\n" % (k, v))
    conf.close()
    log = open(log_path, "w") if log_path else subprocess.DEVNULL

    def _die_with_parent():
        # the master belongs to this job: if the rank that started it dies, the master must not stay behind holding
        # the port (it would otherwise wait for WORKER_FINISH_WORK messages that never come)
        try:
            C.CDLL("libc.so.6", use_errno=True).prctl(1, 15)       # PR_SET_PDEATHSIG, SIGTERM
        except OSError:
            pass
    return subprocess.Popen([exe, "-config", conf.name], stdout=log, stderr=subprocess.STDOUT, preexec_fn=_die_with_parent)


def init(kind="master", master_addr=None, frag_num=1024, world=None, rank=None):
    """Set up the job's group from the launcher's environment (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT, as set by
    ``torchrun`` or ``tools/launch_gpu_job.sh``) and make it current.

    kind = "master": rank 0 starts ``swift_master`` next to itself unless ``master_addr`` (or $SSB200_MASTER_ADDR)
    names one that is already running; every rank registers with it. The master listens on MASTER_PORT + 17 (the
    launcher's own rendezvous store owns MASTER_PORT). The rank the master assigns is the one used from here on; it
    need not equal $RANK (registration order decides), so per-GPU settings must key on LOCAL_RANK, not on the rank.
    kind = "torch": initialise / adopt torch.distributed (NCCL)."""
    env_world = int(os.environ.get("WORLD_SIZE", "1")) if world is None else int(world)
    env_rank = int(os.environ.get("RANK", "0")) if rank is None else int(rank)
    if env_world <= 1:
        return set_current(Group())
    if kind == "torch":
        import torch
        import torch.distributed as dist

        if not dist.is_initialized():
            dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
            dist.init_process_group("nccl", device_id=dev)
        return set_current(TorchGroup())
    if os.environ.get("SSB200_FORBID_TORCH_DIST") == "1":
        # tests: prove that a master-run job never initialises torch.distributed / NCCL
        import torch.distributed as dist

        def _forbidden(*a, **k):
            raise RuntimeError("torch.distributed.init_process_group called in a master-run job")
        dist.init_process_group = _forbidden
    proc = None
    addr = master_addr or os.environ.get("SSB200_MASTER_ADDR")
    if not addr:
        host = os.environ.get("MASTER_ADDR", "127.0.0.1")
        port = int(os.environ.get("MASTER_PORT", "29500")) + 17
        addr = "tcp://%s:%d" % (host, port)
        if env_rank == 0:
            proc = spawn_master(addr, env_world, frag_num, log_path=os.environ.get("SSB200_MASTER_LOG"))
            time.sleep(0.2)
    host = addr.split("//", 1)[1].rsplit(":", 1)[0]
    if host in ("127.0.0.1", "localhost"):
        os.environ.setdefault("SSB200_LOCAL_IP", "127.0.0.1")      # single-box job: listen on loopback (csrc/core/net.h)
    # listen_addr "": an ephemeral port on the local IP (the first non-loopback interface unless SSB200_LOCAL_IP says otherwise)
    return set_current(MasterGroup(addr, listen_addr="", master_proc=proc))
