"""Checkpoint / resume with optimizer state, and the streamed (chunked) dump.

ref: the reference dumps the whole server-side Param through operator<< (parameter/sparsetable.h:49-56,
system/server/init.h:138-149), so an AdaGrad Param carries its accumulator; it never reads a dump back (the loader
is new). Here: "push g1, push g2" must equal "push g1, dump, load into a fresh table, push g2" — rows and
accumulators — through the binary format and through Param-style text."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _keys(n, seed):
    return np.unique(np.random.default_rng(seed).integers(0, 1 << 30, n).astype(np.uint32))


def _dev(a, device):
    from swiftsnails_b200.ops._util import u32_tensor

    return u32_tensor(a, device)


def _state_of(table, keys_dev):
    """AdaGrad accumulator rows of the given keys (via the chunked dump, matched by key)."""
    acc = {}
    for sh in table.owned:
        for k, _, st in table.iter_shard_chunks(sh, with_state=True):
            for kk, ss in zip(k, st):
                acc[int(kk)] = ss
    return acc


@pytest.mark.parametrize("fmt", ["binary", "text"])
def test_resume_continues_adagrad_exactly(device, tmp_path, fmt):
    from swiftsnails_b200.ops.table import ShardedTable
    from swiftsnails_b200.utils import dump

    dim = 16
    keys = _keys(3000, 11)
    kd = _dev(keys, device)
    torch.manual_seed(3)
    g1 = torch.randn(len(keys), dim, device=device)
    g2 = torch.randn(len(keys), dim, device=device)

    a = ShardedTable.local(4096, dim, device, num_shards=2, frag_num=8, with_state=True, seed=77)
    _, sa = a.pull(kd)
    a.push_adagrad(sa, g1.clone(), lr=0.05, eps=1e-8, zero_grads=False)
    a.push_adagrad(sa, g2.clone(), lr=0.05, eps=1e-8, zero_grads=False)
    want, _ = a.pull(kd, insert=False)

    b = ShardedTable.local(4096, dim, device, num_shards=2, frag_num=8, with_state=True, seed=77)
    _, sb = b.pull(kd)
    b.push_adagrad(sb, g1.clone(), lr=0.05, eps=1e-8, zero_grads=False)
    path = str(tmp_path / ("ckpt." + fmt))
    if fmt == "binary":
        dump.save_table_binary(b, path, max_bytes=64 << 10)            # small pieces: many chunks per shard
    else:
        assert dump.save_table(b, path, with_state=True, digits=9, max_bytes=64 << 10) == len(keys)
        assert "\tAcc:\t" in open(path).readline()

    c = ShardedTable.local(8192, dim, device, num_shards=3, frag_num=9, with_state=True, seed=5)   # different sharding
    n = (dump.load_table_binary if fmt == "binary" else dump.load_table)(c, path)
    assert n == len(keys)
    _, sc = c.pull(kd, insert=False)
    c.push_adagrad(sc, g2.clone(), lr=0.05, eps=1e-8, zero_grads=False)
    got, _ = c.pull(kd, insert=False)
    assert torch.allclose(got, want, rtol=1e-6, atol=1e-7)
    acc_a, acc_c = _state_of(a, kd), _state_of(c, kd)
    assert len(acc_c) == len(keys)
    for k in keys[:200]:
        assert np.allclose(acc_a[int(k)], acc_c[int(k)], rtol=1e-6, atol=1e-9)

    # without the state the second update takes a different (larger) step: the state is what makes the resume exact
    d = ShardedTable.local(8192, dim, device, num_shards=3, frag_num=9, with_state=True, seed=5)
    rows_only = str(tmp_path / "rows_only.txt")
    dump.save_table(b, rows_only, digits=9)
    dump.load_table(d, rows_only)
    _, sd = d.pull(kd, insert=False)
    d.push_adagrad(sd, g2.clone(), lr=0.05, eps=1e-8, zero_grads=False)
    off, _ = d.pull(kd, insert=False)
    assert not torch.allclose(off, want, rtol=1e-3, atol=1e-5)


def test_chunked_dump_equals_whole_shard_dump(device):
    from swiftsnails_b200.ops.table import ShardedTable

    t = ShardedTable.local(4096, 8, device, num_shards=1, with_state=True)
    keys = _keys(2500, 4)
    kd = _dev(keys, device)
    rows, gs = t.pull(kd)
    t.push_adagrad(gs, torch.randn_like(rows), lr=0.1, eps=1e-8, zero_grads=False)
    sh = t.owned[0]
    k0, r0 = t.dump_shard(sh)
    whole = {int(k): r for k, r in zip(k0, r0)}
    pieces, n_chunks = {}, 0
    for k, r, st in t.iter_shard_chunks(sh, with_state=True, max_bytes=16 << 10):
        n_chunks += 1
        assert st is not None and st.shape == r.shape and (st >= 0).all()
        for kk, rr in zip(k, r):
            assert int(kk) not in pieces
            pieces[int(kk)] = rr
    assert n_chunks > 4 and len(pieces) == len(whole) == len(keys)
    for k in keys:
        assert np.array_equal(pieces[int(k)], whole[int(k)])
