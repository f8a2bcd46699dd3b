"""The multi-GPU push path with several ranks EMULATED on one GPU, so that the driver's one-GPU test box exercises it
on every run: routing to the owner, private inbox sub-regions, bf16 rows on the wire, count publication and hand-off
flags, parity double-buffering, the owner's segmented apply with a named rule, the bf16 row mirror after many writers.

Kernels that wait for one another must not run as separate concurrent launches on one device, so the virtual ranks
share one process and one stream: every rank's `ship` half runs first, then every rank's `apply` half, whose hand-off
flags are therefore already raised (the flag-wait kernel finds them set; nothing ever waits for another launch). The
real cross-GPU run is tests/test_multigpu.py (2 GPUs; logs for 4 and 8 under profiles/).
Also: VMM allocation -> POSIX fd export -> import -> second mapping, inside one process (the calls that carry a shard
to a peer, parallel/peer.py).
ref: global_push_access.h:123-149 + system/server/init.h:116-132 ("send (key, grad)*, the server applies")."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bf16(x):
    return torch.from_numpy(np.ascontiguousarray(x)).to(torch.bfloat16).to(torch.float32).numpy()


def _setup(device, R, dim, cap_rows, inbox_rows, state_words, bf16_rows, shadow=False, init_kind=0, owner_pulls=None):
    from swiftsnails_b200.ops.table import ShardedTable
    from swiftsnails_b200.parallel.cluster import Inbox

    table = ShardedTable.local(cap_rows, dim, device, num_shards=R, frag_num=64, with_state=state_words, with_shadow=shadow,
                               seed=9, init_kind=init_kind)
    nbytes = Inbox.segment_bytes(inbox_rows, dim, R, bf16_rows)
    segs = [torch.zeros(nbytes, dtype=torch.uint8, device=device) for _ in range(R)]
    ptrs = [t.data_ptr() for t in segs]
    inboxes = [Inbox(table, inbox_rows, device, bf16_rows=bf16_rows, world=R, rank=r, segment_ptrs=ptrs,
                     local_shard=table.owned[r], owner_pulls=owner_pulls) for r in range(R)]
    return table, inboxes, segs


@pytest.mark.parametrize("R,bf16_rows,owner_pulls", [(2, True, True), (4, True, True), (4, True, False), (4, False, False)])
def test_sgd_keys_shared_between_senders(device, R, bf16_rows, owner_pulls):
    """SGD is order-independent: w = w0 - lr * (sum of the gradients), where a gradient that crossed to another owner
    as bf16 was rounded on the wire and one applied by its own owner stayed fp32. Three steps = both parities."""
    from swiftsnails_b200.ops._util import to_u32_numpy, u32_tensor

    dim, lr = 64, 0.5
    table, inboxes, _ = _setup(device, R, dim, 8192, 16384, 0, bf16_rows, owner_pulls=owner_pulls)
    assert all(ib.owner_pulls == bool(owner_pulls) for ib in inboxes)   # outbox + owner reads, or inbox + owner applies local rows
    rng = np.random.default_rng(0)
    pool = np.unique(rng.integers(0, 1 << 30, 6000).astype(np.uint32))
    w0, _ = table.pull(u32_tensor(pool, device))
    want = {int(k): w0[i].cpu().numpy().astype(np.float64) for i, k in enumerate(pool)}
    for step in range(3):
        ctxs = []
        for r in range(R):
            keys = rng.choice(pool, 3000, replace=False)
            _, gslot = table.pull(u32_tensor(keys, device), insert=False)
            g = torch.randn(len(keys), dim, device=device)
            owner = to_u32_numpy(gslot) // table.capacity
            gn = g.cpu().numpy()
            gq = _bf16(gn) if bf16_rows else gn
            for i, k in enumerate(keys):
                want[int(k)] -= lr * (gn[i] if owner[i] == r else gq[i]).astype(np.float64)
            ctxs.append(inboxes[r].ship(gslot, g, lr, adagrad=False))
        for r in range(R):
            inboxes[r].apply(ctxs[r])
    torch.cuda.synchronize()
    got, _ = table.pull(u32_tensor(pool, device), insert=False)
    assert np.allclose(got.cpu().numpy(), np.stack([want[int(k)] for k in pool]), rtol=1e-5, atol=1e-5)
    for ib in inboxes:
        assert int(ib.overflow.item()) == 0 and int(ib.err.item()) == 0
        assert ib.flags_l.tolist() == [3] * R                                   # every sender's hand-off flag carries step 3


@pytest.mark.parametrize("path", ["owner_pulls", "inbox", "inbox_tracked"])
@pytest.mark.parametrize("rule,words,hy", [("adagrad", 1, dict(lr=0.05, eps=1e-8)), ("ftrl", 2, dict(lr=0.1, eps=1.0, a=0.01, b=0.0))])
def test_named_rule_over_four_emulated_ranks_with_mirror(device, rule, words, hy, path):
    """Disjoint keys per sender -> an exact oracle for stateful rules; the table keeps a bf16 mirror, which must be
    bf16(row) bit for bit after own-row updates (single writer) and inbox applies (trailing refresh)."""
    from swiftsnails_b200.ops import rules
    from swiftsnails_b200.ops._util import to_u32_numpy, u32_tensor
    from swiftsnails_b200.utils.reference import adagrad_ref, ftrl_ref

    R, dim = 4, 128
    rl = rules.Rule.get(rule)
    table, inboxes, _ = _setup(device, R, dim, 8192, 16384, words, True, shadow=True, init_kind=rl.init_kind,
                               owner_pulls=(path == "owner_pulls"))
    if path == "inbox_tracked":                  # second-writer tracking: only multi-writer rows are refreshed
        assert all(ib.enable_tracking() for ib in inboxes)
    rng = np.random.default_rng(1)
    pool = np.unique(rng.integers(0, 1 << 30, 8000).astype(np.uint32))
    parts = np.array_split(rng.permutation(pool), R)
    w0 = {r: table.pull(u32_tensor(parts[r], device))[0].cpu().numpy().astype(np.float64) for r in range(R)}
    st = {r: [np.zeros_like(w0[r]), np.zeros_like(w0[r])] for r in range(R)}
    for step in range(3):
        ctxs = []
        for r in range(R):
            kd = u32_tensor(parts[r], device)
            _, gslot = table.pull(kd, insert=False)
            g = (torch.randn(len(parts[r]), dim, device=device) * 0.1)
            owner = to_u32_numpy(gslot) // table.capacity
            gn = g.cpu().numpy()
            g_eff = np.where((owner == r)[:, None], gn, _bf16(gn)).astype(np.float64)
            if rule == "adagrad":
                st[r][0] = st[r][0] + g_eff * g_eff
                w0[r] = w0[r] - hy["lr"] * g_eff / np.sqrt(st[r][0] + hy["eps"])
            else:
                w0[r], st[r][0], st[r][1] = ftrl_ref(w0[r], st[r][0], st[r][1], g_eff, hy["lr"], hy["eps"], hy["a"], hy["b"])
            ctxs.append(inboxes[r].ship(gslot, g, hy["lr"], hy["eps"], rule=rule, **{k: v for k, v in hy.items() if k in "abc"}))
        for r in range(R):
            inboxes[r].apply(ctxs[r])
    torch.cuda.synchronize()
    for r in range(R):
        kd = u32_tensor(parts[r], device)
        got, gsl = table.pull(kd, insert=False)
        assert 0 < int((to_u32_numpy(gsl) // table.capacity != r).sum()) < len(parts[r])          # some rows really crossed
        assert np.allclose(got.cpu().numpy(), w0[r], rtol=3e-4, atol=3e-6)
        bf = torch.empty(len(parts[r]), dim, dtype=torch.bfloat16, device=device)
        table.pull(kd, rows_bf16=bf, want_fp32=False, insert=False)
        assert torch.equal(bf, got.to(torch.bfloat16))
    assert all(int(ib.err.item()) == 0 and int(ib.overflow.item()) == 0 for ib in inboxes)


def test_overflow_is_counted_and_a_silent_sender_is_reported_not_trapped(device):
    """A sub-region that is too small drops rows and counts them; a sender that never ships leaves its flag low: the
    bounded wait gives up, sets err (the trainers read it and fail) and the process lives on."""
    from swiftsnails_b200.ops._util import u32_tensor

    R, dim = 2, 64
    table, inboxes, _ = _setup(device, R, dim, 8192, 512, 0, True)              # 256 rows per (owner, sender)
    keys = np.unique(np.random.default_rng(2).integers(0, 1 << 30, 4000).astype(np.uint32))
    _, gslot = table.pull(u32_tensor(keys, device))
    g = torch.randn(len(keys), dim, device=device)
    inboxes[0].wait_limit = 2000                                                # iterations: give up quickly
    c0 = inboxes[0].ship(gslot, g, 0.1, adagrad=False)
    inboxes[0].apply(c0)                                                        # rank 1 never shipped this step
    torch.cuda.synchronize()
    assert int(inboxes[0].overflow.item()) > 0
    assert int(inboxes[0].err.item()) != 0


def test_vmm_export_import_second_mapping_in_one_process(device):
    """cuMemCreate -> export as a POSIX fd -> import -> map at a second address: a write through one mapping is read
    through the other (what parallel/peer.py does between processes, over SCM_RIGHTS)."""
    from swiftsnails_b200 import _lib
    from swiftsnails_b200.parallel.peer import _vmm, tensor_from_ptr, vmm_granularity

    dev_index = device.index if device.index is not None else torch.cuda.current_device()
    gran = vmm_granularity(dev_index)
    size = 2 * gran
    handle, fd = C.c_uint64(), C.c_int()
    _vmm("ssb_vmm_create", C.c_int(dev_index), C.c_uint64(size), C.byref(handle), C.byref(fd))
    va1, va2 = C.c_uint64(), C.c_uint64()
    _vmm("ssb_vmm_reserve", C.c_uint64(size), C.c_uint64(gran), C.byref(va1))
    _vmm("ssb_vmm_map", C.c_uint64(va1.value), C.c_uint64(size), C.c_uint64(handle.value), C.c_int(dev_index))
    h2 = C.c_uint64()
    _vmm("ssb_vmm_import", C.c_int(fd.value), C.byref(h2))
    os.close(fd.value)
    _vmm("ssb_vmm_reserve", C.c_uint64(size), C.c_uint64(gran), C.byref(va2))
    _vmm("ssb_vmm_map", C.c_uint64(va2.value), C.c_uint64(size), C.c_uint64(h2.value), C.c_int(dev_index))
    assert va1.value != va2.value
    a = tensor_from_ptr(va1.value, size, device)
    b = tensor_from_ptr(va2.value, size, device)
    a.copy_(torch.arange(size, device=device, dtype=torch.int64).to(torch.uint8))
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    b[:4096].zero_()
    torch.cuda.synchronize()
    assert int(a[:4096].sum().item()) == 0 and int(a[4096:8192].to(torch.int64).sum().item()) > 0
