"""Build the sharded table for a job: one server shard per GPU, every shard visible to every
worker.  Single process => plain local shard; several ranks (``parallel.group``: registered with the
master over ZeroMQ, or a torch.distributed job) => VMM peer mapping over NVLink (``peer.PeerMemory``);
the fragment map is the one the master built (NODE_ASKFOR_HASHFRAG) when the master runs the job.

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
    shadow_off = state_off + rows_bytes * int(with_state)       # with_state = optimizer-state words per weight (True = 1)
    total = shadow_off + (rows_bytes // 2 if with_shadow else 0)
    return {"keys_off": 0, "rows_off": rows_off, "state_off": state_off if with_state else None,
            "shadow_off": shadow_off if with_shadow else None, "total": total, "pitch": pitch}


def make_table(rows_per_shard, dim, device, with_state=False, init_scale=None, frag_num=1024, seed=0x5EED,
               load_factor=0.5, force_peer=False, with_shadow=False, init_kind=0):
    from . import group as _group

    grp = _group.current()
    world, rank = grp.world, grp.rank
    capacity = round_capacity(rows_per_shard, load_factor)
    hf = grp.hashfrag(frag_num) if world > 1 else HashFrag(world, max(frag_num, world))
    if world == 1 and not force_peer:
        shard = DeviceShard(capacity, dim, device, with_state, with_shadow=with_shadow)
        return ShardedTable(hf, capacity, dim, [shard.ptrs()], device, seed=seed, init_scale=init_scale,
                            owned=[shard], rank=0, init_kind=init_kind)
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
        grp.barrier()  # every shard is initialised before anyone probes it
    ptrs = []
    for r in range(world):
        b = pm.segment_ptr(r)
        ptrs.append((b + lay["keys_off"], b + lay["rows_off"], (b + lay["state_off"]) if with_state else 0,
                     (b + lay["shadow_off"]) if with_shadow else 0))
    table = ShardedTable(hf, capacity, dim, ptrs, device, seed=seed, init_scale=init_scale, owned=[shard],
                         rank=rank, init_kind=init_kind)
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

    @staticmethod
    def segment_bytes(capacity, dim, world, bf16_rows):
        """Bytes of one rank's inbox segment (two parities + flags) for these parameters."""
        seg_rows = max(64, (int(capacity) + world - 1) // world)
        cap = seg_rows * world
        count_off = cap * dim * (2 if bf16_rows else 4) + cap * 4
        per = (count_off + 256 + 4095) // 4096 * 4096
        return 2 * per + 4096

    def __init__(self, table: ShardedTable, capacity: int, device, bf16_rows: bool = False, barrier: str = "flags",
                 world=None, rank=None, segment_ptrs=None, local_shard=None, owner_pulls=None):
        """``world`` / ``rank`` / ``segment_ptrs`` / ``local_shard``: for callers that bring their own memory — tests
        that emulate several ranks inside one process on one GPU (every rank's segment is then a plain device buffer
        of ``segment_bytes`` bytes, zeroed) — instead of the job's group and a :class:`PeerMemory` exchange."""
        import ctypes as C

        from . import group as _group
        from .peer import tensor_from_ptr

        self.table, self.device = table, device
        self.group = _group.current() if world is None else None
        self.world, self.rank = (self.group.world, self.group.rank) if world is None else (int(world), int(rank))
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
        if segment_ptrs is None:
            self.pm = PeerMemory(flags_off + 4096, device)
            segment_ptrs = [self.pm.segment_ptr(r) for r in range(self.world)]
            seg = self.pm.local_tensor()
            seg.zero_()
        else:
            seg = tensor_from_ptr(segment_ptrs[self.rank], flags_off + 4096, device)
        # owner_pulls: a sender compacts the rows for each owner into ITS OWN segment (an outbox) and the owner's apply
        # kernel reads them over NVLink while applying them — the shipment pass disappears and the rows cross the
        # fabric as loads. Off by default: measured +3.5 % at 2 GPUs (0.592 vs 0.614 ms/step) but -30 % at 8 (0.943
        # vs 0.728): the pull of the next step already fills the INBOUND direction of every GPU's links, and the
        # sender-side stores use the otherwise idle outbound direction. Default (False): senders store into the
        # owners' segments (inboxes) and owners apply local memory.
        self.owner_pulls = bool(owner_pulls) and self.bf16_rows
        self.views, self.src_views, self.rows_l, self.slots_l, self.count_l = [], [], [], [], []
        for par in range(2):
            # the view a SENDER uses: one private sub-region per owner — on the owner (inbox) or here (outbox)
            v = _lib.SsbInboxView()
            for r in range(self.world):
                if self.owner_pulls:
                    b = segment_ptrs[self.rank] + par * per
                    v.rows[r] = b + r * self.seg_rows * D * esz
                    v.slots[r] = b + slots_off + r * self.seg_rows * 4
                    v.count[r] = b + count_off + r * 4
                else:
                    b = segment_ptrs[r] + par * per
                    v.rows[r] = b + self.rank * self.seg_rows * D * esz
                    v.slots[r] = b + slots_off + self.rank * self.seg_rows * 4
                    v.count[r] = b + count_off + self.rank * 4
                v.flags[r] = segment_ptrs[r] + flags_off
            v.capacity = self.seg_rows
            self.views.append(v)
            # what an OWNER reads in pull mode: sender y's sub-region for this rank, in y's segment
            sv = _lib.SsbInboxView()
            for y in range(self.world):
                b = segment_ptrs[y] + par * per
                sv.rows[y] = b + self.rank * self.seg_rows * D * esz
                sv.slots[y] = b + slots_off + self.rank * self.seg_rows * 4
                sv.count[y] = b + count_off + self.rank * 4
                sv.flags[y] = segment_ptrs[y] + flags_off
            sv.capacity = self.seg_rows
            self.src_views.append(sv)
            # what the OWNER applies: all sub-regions of its own inbox
            base = par * per
            rows = seg[base:base + rows_b]
            self.rows_l.append(rows.view(torch.bfloat16 if self.bf16_rows else torch.float32).view(self.capacity, D))
            self.slots_l.append(seg[base + slots_off:base + slots_off + self.capacity * 4].view(torch.int32))
            self.count_l.append(seg[base + count_off:base + count_off + 4 * self.world].view(torch.int32))
        self.flags_l = seg[flags_off:flags_off + 4 * self.world].view(torch.int32)
        shard = local_shard if local_shard is not None else table.owned[0]
        self.local = ShardedTable(HashFrag(1, 1), table.capacity, D, [shard.ptrs()], device, seed=table.seed,
                                  init_scale=table.init_scale, owned=[shard], init_kind=table.init_kind)
        self.overflow = torch.zeros(1, dtype=torch.int32, device=device)
        self.err = torch.zeros(1, dtype=torch.int32, device=device)         # a hand-off wait ran out (a sender never showed up)
        self.cursor = torch.zeros(16, dtype=torch.int32, device=device)     # rows claimed per owner this step
        self.token = torch.zeros(1, dtype=torch.int32, device=device)
        self.parity = 0
        self.step = 0
        self.ship_first = True       # see push()
        self.wait_limit = 0          # iterations (~0.2 us each) the hand-off wait may spin; 0 = the library default (~27 s)
        self.own_stream = torch.cuda.Stream(device=device)      # own-row updates run next to the shipment (None: in line)
        self.own_done = torch.cuda.Event()
        # Second-writer tracking of the inbox apply (ssb::MirrorTrack: only rows that several senders carry get the
        # trailing mirror refresh) is available through enable_tracking() but off: with the tracked apply every
        # writer needs the atomic's return value and stores the mirror itself, and about half of the incoming rows
        # have a second writer anyway — measured 132 vs 119 us (2 ranks) and 246 vs 201 us (8 ranks) per apply of
        # 152 K / 266 K rows (benchmarks/ship_probe.py). None = refresh every applied row.
        self.track = None
        self._C = C
        torch.cuda.synchronize(device)
        if self.group is not None:
            self.group.barrier()

    def enable_tracking(self):
        shard = self.local.owned[0]
        if shard.shadow is not None and self.bf16_rows:
            self.track = (torch.zeros(self.table.capacity, dtype=torch.int32, device=self.device),)
        return self.track is not None

    def push(self, gslot, grads, lr, eps=1e-8, adagrad=True, n_dev=None, n=None, rule=None, **hyper):
        """``rule``: a named device-side update rule (ops.rules) instead of the built-in SGD / AdaGrad: the send
        kernel then only ships, this rank's own rows and the inbox are applied by the rule's own entry points."""
        ctx = self.ship(gslot, grads, lr, eps, adagrad, n_dev, n, rule, **hyper)
        self.apply(ctx)

    def ship(self, gslot, grads, lr, eps=1e-8, adagrad=True, n_dev=None, n=None, rule=None, **hyper):
        """First half of :meth:`push`: ship the rows other ranks own, raise the hand-off flags, apply my own rows.
        Returns the context for :meth:`apply`. (Split so that a test can run every emulated rank's ship before any
        rank's apply on one stream — no kernel then ever waits for another launch.)"""
        C = self._C
        n = gslot.numel() if n is None else n
        par = self.parity
        self.parity ^= 1
        self.step += 1
        sp = _lib.stream_ptr()
        view = C.byref(self.views[par])
        marks = getattr(self, "marks", None)       # benchmarks: a list that receives (name, CUDA event) per stage

        def mark(name):
            if marks is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                marks.append((name, e))
        mark("start")
        from ..ops import rules as _rules

        # Ship first: the send kernel only ships the rows other ranks own and the hand-off flags go up right behind
        # it; this rank's own rows are applied AFTERWARDS, while the peers' shipments are still landing, so the wait
        # that follows finds its flags already raised. (ship_first = False: the older send kernel that applies the
        # own rows between its shipments — built-in SGD / AdaGrad only — kept for comparison.)
        ship_first = getattr(self, "ship_first", True) or rule is not None or self.owner_pulls
        if ship_first:
            rule = rule if rule is not None else ("adagrad" if adagrad else "sgd")
            rule = rule if isinstance(rule, _rules.Rule) else _rules.Rule.get(rule)
            hy = rule.hyper(lr=lr, eps=eps, **hyper)
        if ship_first and self.own_stream is not None:
            ev_ready = torch.cuda.Event()
            ev_ready.record()              # the gradients are complete here (stream order): the side stream may start
        _lib.call("ssb_push_send2", self.table._view_ref(), view, gslot.data_ptr(), grads.data_ptr(), n,
                  _lib.ptr(n_dev), self.rank, float(lr), float(eps), -1 if ship_first else (1 if adagrad else 0),
                  1 if self.bf16_rows else 0, self.overflow.data_ptr(), self.cursor.data_ptr(), sp)
        mark("send")
        if self.barrier == "flags":
            _lib.call("ssb_inbox_publish", view, self.world, self.rank, self.cursor.data_ptr(), self.step, sp)
        else:
            _lib.call("ssb_inbox_publish", view, self.world, self.rank, self.cursor.data_ptr(), 0, sp)
        if ship_first:         # my own rows
            if self.own_stream is not None:
                # next to the shipment, on a side stream: shipping is bound by NVLink, this update by HBM atomics
                cur = torch.cuda.current_stream()
                self.own_stream.wait_event(ev_ready)
                with torch.cuda.stream(self.own_stream):
                    rule.push(self.table, gslot, grads, hy, n_dev=n_dev, zero_grads=False, n=n, only_shard=self.rank)
                    self.own_done.record()
                gslot.record_stream(self.own_stream)
                grads.record_stream(self.own_stream)
            else:
                rule.push(self.table, gslot, grads, hy, n_dev=n_dev, zero_grads=False, n=n, only_shard=self.rank)
        mark("own")
        return dict(par=par, step=self.step, ship_first=ship_first, rule=rule if ship_first else None, hy=hy if ship_first else None,
                    lr=lr, eps=eps, adagrad=adagrad, mark=mark)

    def apply(self, ctx):
        """Second half of :meth:`push`: wait for every sender's hand-off, apply my inbox to my shard."""
        par, ship_first, rule, hy, lr, eps, adagrad, mark = (ctx[k] for k in ("par", "ship_first", "rule", "hy", "lr", "eps", "adagrad", "mark"))
        sp = _lib.stream_ptr()
        if self.barrier == "flags":
            _lib.call("ssb_flags_wait2", self.flags_l.data_ptr(), self.world, ctx["step"], int(self.wait_limit), self.err.data_ptr(), sp)
        else:
            import torch.distributed as dist         # barrier="nccl": the collective variant needs a torch.distributed job

            dist.all_reduce(self.token)      # every rank's shipment has landed in my inbox
        if ship_first and self.own_stream is not None:
            # own rows and inbox rows can be the same rows: the inbox apply (and its mirror refresh) goes after both
            torch.cuda.current_stream().wait_event(self.own_done)
        mark("handoff")
        if self.owner_pulls and ship_first:
            rule.apply_outbox(self.local, self.src_views[par], self.rank, self.world, self.seg_rows, hy)
        elif self.bf16_rows and ship_first:
            rule.push_bf16_seg(self.local, self.slots_l[par], self.rows_l[par], self.seg_rows, self.count_l[par], self.world, hy,
                               track=None if self.track is None else self.track + (ctx["step"],))
        elif self.bf16_rows:
            _lib.call("ssb_push_bf16_seg", self.local._view_ref(), self.slots_l[par].data_ptr(), self.rows_l[par].data_ptr(),
                      self.seg_rows, self.world, self.count_l[par].data_ptr(), float(lr), float(eps), 1 if adagrad else 0, sp)
        else:
            S = self.seg_rows
            for r in range(self.world):
                if r == self.rank:
                    continue                 # own rows were applied directly by the send kernel
                sl, rw, cn = self.slots_l[par][r * S:(r + 1) * S], self.rows_l[par][r * S:(r + 1) * S], self.count_l[par][r:r + 1]
                if ship_first:
                    rule.push(self.local, sl, rw, hy, n_dev=cn, zero_grads=False, n=S)
                elif adagrad:
                    self.local.push_adagrad(sl, rw, lr, eps, n_dev=cn, zero_grads=False, n=S)
                else:
                    self.local.push_sgd(sl, rw, lr, n_dev=cn, zero_grads=False, n=S)
        mark("apply")
        # no clearing: every sender overwrites its count on its next publish (also with 0 rows)


class ChunkInbox:
    """Chunked owner-side inboxes: shipping and applying overlap (csrc/kernels/table.cu, push_send_chunks_kernel /
    push_apply_chunks_kernel). Same idea as :class:`Inbox` — every rank owns two inboxes (step parity) in
    peer-mapped memory, one private sub-region per sender, gradient rows as bf16 on the wire — but a sub-region is
    cut into 256-row chunks: the send kernel raises a chunk's ready flag on the owner the moment the chunk is
    complete and publishes the total at its end; the owner's apply kernel runs on a second stream NEXT TO its own
    send kernel and applies (sender, chunk) items as they land. Flags carry the step number, nothing is cleared.
    ref: global_push_access.h:123-149, system/server/init.h:116-132 ("send (key, grad)*, the server applies").

    Status: opt-in (``attach_inbox(..., barrier="chunks")``, ``bench.py --inbox-barrier chunks``). Correct on 2 GPUs
    (no lost rows, same loss) but measured SLOWER than the flag hand-off there, 0.83 vs 0.59 ms/step: the apply
    kernel is held to one block per SM so that its waiting blocks can never crowd the send kernel (and the rest
    of the pipelined step) off the SMs, and one block per SM does not keep enough row updates in flight; a
    variant with more blocks / a leaner chunk walk ran into its wait bound on 2 GPUs and was dropped.

    ``segment_ptrs``: base device pointers of every rank's inbox segment, for callers that bring their own
    (peer-mapped or, in single-process tests that emulate several ranks on one GPU, plain) memory; default: a
    :class:`PeerMemory` exchange over ``torch.distributed``."""

    CHUNK = 256

    @staticmethod
    def layout(capacity, dim, world):
        seg_rows = max(ChunkInbox.CHUNK, (int(capacity) + world - 1) // world)
        seg_rows = (seg_rows + ChunkInbox.CHUNK - 1) // ChunkInbox.CHUNK * ChunkInbox.CHUNK
        cap = seg_rows * world
        max_chunks = seg_rows // ChunkInbox.CHUNK
        rows_b = cap * dim * 2
        slots_off = rows_b
        ready_off = slots_off + cap * 4
        total_off = ready_off + world * max_chunks * 4
        per = (total_off + world * 8 + 4095) // 4096 * 4096
        return {"seg_rows": seg_rows, "capacity": cap, "max_chunks": max_chunks, "slots_off": slots_off,
                "ready_off": ready_off, "total_off": total_off, "per": per, "total": 2 * per}

    def __init__(self, table: ShardedTable, capacity: int, device, world=None, rank=None, segment_ptrs=None, local_shard=None,
                 spin_limit=0):
        import ctypes as C

        from .peer import tensor_from_ptr

        self.table, self.device = table, device
        if world is None:
            from . import group as _group

            world, rank = _group.current().world, _group.current().rank
        self.world, self.rank = int(world), int(rank)
        D = table.dim
        lay = self.layout(capacity, D, self.world)
        self.lay = lay
        self.seg_rows, self.capacity, self.max_chunks = lay["seg_rows"], lay["capacity"], lay["max_chunks"]
        if segment_ptrs is None:
            self.pm = PeerMemory(lay["total"], device)
            segment_ptrs = [self.pm.segment_ptr(r) for r in range(self.world)]
            seg = self.pm.local_tensor()
            seg.zero_()
        else:
            seg = tensor_from_ptr(segment_ptrs[self.rank], lay["total"], device)
        self.views, self.rows_l, self.slots_l, self.ready_l, self.total_l = [], [], [], [], []
        for par in range(2):
            v = _lib.SsbChunkInboxView()
            for r in range(self.world):
                b = segment_ptrs[r] + par * lay["per"]
                v.rows[r] = b + self.rank * self.seg_rows * D * 2
                v.slots[r] = b + lay["slots_off"] + self.rank * self.seg_rows * 4
                v.ready[r] = b + lay["ready_off"] + self.rank * self.max_chunks * 4
                v.total[r] = b + lay["total_off"] + self.rank * 8
            v.seg_rows, v.max_chunks = self.seg_rows, self.max_chunks
            self.views.append(v)
            base = par * lay["per"]
            self.rows_l.append(seg[base:base + self.capacity * D * 2].view(torch.bfloat16).view(self.capacity, D))
            self.slots_l.append(seg[base + lay["slots_off"]:base + lay["slots_off"] + self.capacity * 4].view(torch.int32))
            self.ready_l.append(seg[base + lay["ready_off"]:base + lay["ready_off"] + self.world * self.max_chunks * 4].view(torch.int32))
            self.total_l.append(seg[base + lay["total_off"]:base + lay["total_off"] + self.world * 8].view(torch.int32))
        shard = local_shard if local_shard is not None else table.owned[0]
        self.local = ShardedTable(HashFrag(1, 1), table.capacity, D, [shard.ptrs()], device, seed=table.seed,
                                  init_scale=table.init_scale, owned=[shard])
        i32 = dict(dtype=torch.int32, device=device)
        self.overflow = torch.zeros(1, **i32)
        self.err = torch.zeros(1, **i32)             # an apply wait ran out (a sender never delivered)
        self.cursor = torch.zeros(16, **i32)
        self.done = torch.zeros(16 * self.max_chunks, **i32)
        self.finished = torch.zeros(1, **i32)
        self.apply_stream = torch.cuda.Stream(device=device)
        self.spin_limit = int(spin_limit)
        self.parity = 0
        self.step = 0
        self._C = C
        torch.cuda.synchronize(device)

    def send(self, gslot, grads, lr, eps=1e-8, adagrad=True, n_dev=None, n=None):
        """Apply my own rows, ship the others (current stream). Returns (parity, tag) for :meth:`apply`."""
        n = gslot.numel() if n is None else n
        par = self.parity
        self.parity ^= 1
        self.step += 1
        _lib.call("ssb_push_send_chunks", self.table._view_ref(), self._C.byref(self.views[par]), gslot.data_ptr(),
                  grads.data_ptr(), n, _lib.ptr(n_dev), self.rank, float(lr), float(eps), 1 if adagrad else 0, self.step,
                  self.overflow.data_ptr(), self.cursor.data_ptr(), self.done.data_ptr(), self.finished.data_ptr(),
                  _lib.stream_ptr())
        return par, self.step

    def apply(self, par, tag, lr, eps=1e-8, adagrad=True):
        """Apply every sender's chunks of my inbox as they land (current stream)."""
        _lib.call("ssb_push_apply_chunks", self.local._view_ref(), self.slots_l[par].data_ptr(), self.rows_l[par].data_ptr(),
                  self.ready_l[par].data_ptr(), self.total_l[par].data_ptr(), self.seg_rows, self.world, self.rank, int(tag),
                  float(lr), float(eps), 1 if adagrad else 0, self.spin_limit, self.err.data_ptr(), _lib.stream_ptr())

    def push(self, gslot, grads, lr, eps=1e-8, adagrad=True, n_dev=None, n=None):
        cur = torch.cuda.current_stream()
        start = torch.cuda.Event()
        start.record(cur)
        par, tag = self.send(gslot, grads, lr, eps, adagrad, n_dev, n)
        self.apply_stream.wait_event(start)           # same dependencies as the send, not the send itself: they overlap
        with torch.cuda.stream(self.apply_stream):
            self.apply(par, tag, lr, eps, adagrad)
        cur.wait_stream(self.apply_stream)            # the next push (and the caller) come after both


def attach_inbox(table: ShardedTable, capacity: int, bf16_rows: bool = False, barrier: str = "flags", owner_pulls=None):
    """Give a peer-mapped table owner-side inboxes (no-op for a single process)."""
    from . import group as _group

    grp = _group.current()
    if grp.world == 1:
        table.inbox = None
        return None
    if barrier == "nccl" and grp.kind != "torch":
        raise ValueError('barrier="nccl" needs a torch.distributed job (parallel.group.init("torch"))')
    if barrier == "chunks":
        table.inbox = ChunkInbox(table, capacity, table.device)
        grp.barrier()
    else:
        table.inbox = Inbox(table, capacity, table.device, bf16_rows=bf16_rows, barrier=barrier, owner_pulls=owner_pulls)
    return table.inbox
