"""Chunked inbox push (ship / apply overlap) with several ranks EMULATED on one GPU.

The driver's GPU-test box has one GPU, and kernels of different "ranks" that wait for one another must not run as
concurrent launches on one device — so here all virtual ranks live in one process on one stream: every rank's send
kernel runs first, then every rank's apply kernel, whose chunk flags / totals are therefore already in place (no
kernel ever waits on another launch). That exercises the whole data path — routing to the owner, private sub-regions,
chunk completion counters, ready flags, totals, parity double-buffering, tag reuse, bf16 rows on the wire, the owner's
apply — against a NumPy oracle. The real cross-GPU run is tests/test_multigpu.py.
ref: global_push_access.h:123-149 + system/server/init.h:116-132 (send (key, grad)*, the server applies)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bf16(x):
    return torch.from_numpy(np.ascontiguousarray(x)).to(torch.bfloat16).to(torch.float32).numpy()


def _setup(device, R, dim, cap_rows, inbox_rows, with_state):
    from swiftsnails_b200.ops.table import ShardedTable
    from swiftsnails_b200.parallel.cluster import ChunkInbox

    table = ShardedTable.local(cap_rows, dim, device, num_shards=R, frag_num=64, with_state=with_state, seed=9)
    lay = ChunkInbox.layout(inbox_rows, dim, R)
    segs = [torch.zeros(lay["total"], dtype=torch.uint8, device=device) for _ in range(R)]
    ptrs = [t.data_ptr() for t in segs]
    inboxes = [ChunkInbox(table, inbox_rows, device, world=R, rank=r, segment_ptrs=ptrs, local_shard=table.owned[r])
               for r in range(R)]
    return table, inboxes, segs


@pytest.mark.parametrize("R", [2, 4])
def test_sgd_shared_keys_three_steps(device, R):
    """Keys shared between senders, SGD (order-independent): w = w0 - lr * sum of the gradients, where a gradient that
    crossed to another owner was rounded to bf16 on the wire and one applied by its own owner stayed fp32."""
    from swiftsnails_b200.ops._util import to_u32_numpy, u32_tensor

    dim, lr = 64, 0.5
    table, inboxes, _ = _setup(device, R, dim, 8192, 16384, with_state=False)
    rng = np.random.default_rng(0)
    pool = np.unique(rng.integers(0, 1 << 30, 6000).astype(np.uint32))
    w0, _ = table.pull(u32_tensor(pool, device))
    want = {int(k): w0[i].cpu().numpy().astype(np.float64) for i, k in enumerate(pool)}
    for step in range(3):
        sends = []
        for r in range(R):
            keys = rng.choice(pool, 3000, replace=False)
            kd = u32_tensor(keys, device)
            _, gslot = table.pull(kd, insert=False)
            g = torch.randn(len(keys), dim, device=device)
            owner = to_u32_numpy(gslot) // table.capacity
            gn = g.cpu().numpy()
            gq = _bf16(gn)
            for i, k in enumerate(keys):
                want[int(k)] -= lr * (gn[i] if owner[i] == r else gq[i]).astype(np.float64)
            sends.append(inboxes[r].send(gslot, g, lr, adagrad=False))
        for r in range(R):
            inboxes[r].apply(*sends[r], lr, adagrad=False)
    torch.cuda.synchronize()
    got, _ = table.pull(u32_tensor(pool, device), insert=False)
    got = got.cpu().numpy()
    ref = np.stack([want[int(k)] for k in pool])
    assert np.allclose(got, ref, rtol=1e-5, atol=1e-5)
    for ib in inboxes:
        assert int(ib.overflow.item()) == 0 and int(ib.err.item()) == 0
        assert int(ib.cursor.sum().item()) == 0 and int(ib.done.sum().item()) == 0 and int(ib.finished.item()) == 0


def test_adagrad_disjoint_keys(device):
    from swiftsnails_b200.ops._util import to_u32_numpy, u32_tensor
    from swiftsnails_b200.utils.reference import adagrad_ref

    R, dim, lr, eps = 4, 32, 0.05, 1e-8
    table, inboxes, _ = _setup(device, R, dim, 8192, 16384, with_state=True)
    rng = np.random.default_rng(1)
    pool = np.unique(rng.integers(0, 1 << 30, 8000).astype(np.uint32))
    parts = np.array_split(rng.permutation(pool), R)
    w = {r: table.pull(u32_tensor(parts[r], device))[0] for r in range(R)}
    acc = {r: torch.zeros_like(w[r]) for r in range(R)}
    for step in range(2):
        sends = []
        for r in range(R):
            kd = u32_tensor(parts[r], device)
            _, gslot = table.pull(kd, insert=False)
            g = torch.randn(len(parts[r]), dim, device=device)
            owner = torch.from_numpy((to_u32_numpy(gslot) // table.capacity).astype(np.int64)).to(device)
            g_eff = torch.where((owner == r)[:, None], g, g.to(torch.bfloat16).to(torch.float32))
            w[r], acc[r] = adagrad_ref(w[r], acc[r], g_eff, lr, eps)
            sends.append(inboxes[r].send(gslot, g, lr, eps, adagrad=True))
        for r in range(R):
            inboxes[r].apply(*sends[r], lr, eps, adagrad=True)
    torch.cuda.synchronize()
    for r in range(R):
        got, _ = table.pull(u32_tensor(parts[r], device), insert=False)
        assert torch.allclose(got, w[r], rtol=1e-4, atol=1e-6)
    assert all(int(ib.err.item()) == 0 and int(ib.overflow.item()) == 0 for ib in inboxes)


def test_overflow_is_counted_and_empty_step_still_publishes(device):
    """A sub-region that is too small drops rows and counts them (the trainers read the counter and fail); a rank with
    nothing to ship still publishes its (zero) totals, so the owners' apply does not wait for it."""
    from swiftsnails_b200.ops._util import u32_tensor

    R, dim = 2, 64
    table, inboxes, _ = _setup(device, R, dim, 8192, 512, with_state=False)      # 256 rows per (owner, sender)
    keys = np.unique(np.random.default_rng(2).integers(0, 1 << 30, 4000).astype(np.uint32))
    _, gslot = table.pull(u32_tensor(keys, device))
    g = torch.randn(len(keys), dim, device=device)
    s0 = inboxes[0].send(gslot, g, 0.1, adagrad=False)
    s1 = inboxes[1].send(gslot[:0], g[:0], 0.1, adagrad=False, n=0)              # nothing to ship
    inboxes[0].apply(*s0, 0.1, adagrad=False)
    inboxes[1].apply(*s1, 0.1, adagrad=False)
    torch.cuda.synchronize()
    assert int(inboxes[0].overflow.item()) > 0
    assert int(inboxes[0].err.item()) == 0 and int(inboxes[1].err.item()) == 0
