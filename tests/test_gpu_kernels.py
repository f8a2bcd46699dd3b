"""GPU numerics tests: every kernel against a NumPy / PyTorch fp32 oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _keys(n, hi=2 ** 31, seed=0, dup=1.0):
    rng = np.random.default_rng(seed)
    k = rng.integers(0, int(hi * dup) if dup < 1 else hi, size=n, dtype=np.int64).astype(np.uint32)
    return k


def _dev(a, device):
    from swiftsnails_b200.ops._util import u32_tensor

    return u32_tensor(a, device)


def test_fmix64(device):
    from swiftsnails_b200.ops import hashing
    from swiftsnails_b200.parallel.hashfrag import fmix64

    x = np.random.default_rng(1).integers(0, 2 ** 63, size=10_000, dtype=np.int64)
    out = hashing.fmix64(torch.from_numpy(x).to(device)).cpu().numpy().view(np.uint64)
    assert np.array_equal(out, fmix64(x.view(np.uint64)))


@pytest.mark.parametrize("num_nodes,frag_num", [(2, 100), (7, 100), (8, 1024)])
def test_hash_route_and_bucket(device, num_nodes, frag_num):
    from swiftsnails_b200.ops import hashing
    from swiftsnails_b200.ops._util import to_u32_numpy
    from swiftsnails_b200.parallel.hashfrag import HashFrag

    hf = HashFrag(num_nodes, frag_num)
    keys = _keys(50_000, seed=2)
    kd = _dev(keys, device)
    fm = hf.device_map(device)
    owner, counts = hashing.hash_route(kd, fm, num_nodes)
    ref = hf.to_shard(keys)
    assert np.array_equal(owner.cpu().numpy(), ref)
    assert np.array_equal(counts.cpu().numpy(), np.bincount(ref, minlength=num_nodes))
    bucketed, perm, counts2, offsets = hashing.bucket_by_owner(kd, fm, num_nodes)
    b, p, off = to_u32_numpy(bucketed), to_u32_numpy(perm), offsets.cpu().numpy()
    assert np.array_equal(keys[p], b)
    assert sorted(p.tolist()) == list(range(len(keys)))
    for o in range(num_nodes):
        assert np.all(hf.to_shard(b[off[o]:off[o + 1]]) == o)


def test_dedup(device):
    from swiftsnails_b200.ops._util import to_u32_numpy
    from swiftsnails_b200.ops.dedup import Deduper

    d = Deduper(200_000, device)
    for it in range(3):
        keys = _keys(150_000 + it, hi=40_000, seed=10 + it)
        keys[::1000] = 0xFFFFFFFF  # padding keys are skipped
        unique, inverse, n_unique = d(_dev(keys, device))
        n = int(n_unique.item())
        u, inv = to_u32_numpy(unique)[:n], to_u32_numpy(inverse)
        valid = keys != 0xFFFFFFFF
        assert n == len(np.unique(keys[valid]))
        assert len(np.unique(u)) == n
        assert np.array_equal(u[inv[valid]], keys[valid])
        assert np.all(inv[~valid] == 0xFFFFFFFF)


@pytest.mark.parametrize("dim,num_shards", [(128, 1), (64, 4), (1, 2), (20, 3)])
def test_pull_insert_init_and_gather(device, dim, num_shards):
    from swiftsnails_b200.ops._util import to_u32_numpy
    from swiftsnails_b200.ops.table import ShardedTable
    from swiftsnails_b200.utils.reference import init_rows

    t = ShardedTable.local(4096, dim, device, num_shards=num_shards, frag_num=64, seed=99, init_scale=0.5)
    keys = np.unique(_keys(3000, seed=3))
    kd = _dev(keys, device)
    rows, gslot = t.pull(kd)
    torch.cuda.synchronize()
    ref = init_rows(keys, dim, 99, 0.5)
    assert np.array_equal(rows.cpu().numpy(), ref)
    g = to_u32_numpy(gslot)
    assert len(np.unique(g)) == len(keys)  # distinct slots
    assert np.array_equal(g // t.capacity, t.hashfrag.to_shard(keys))
    assert int(t.stats[0].item()) == len(keys) and int(t.stats[1].item()) == 0
    assert t.size() == len(keys)
    # second pull: same slots, same rows, no new inserts
    rows2, gslot2 = t.pull(kd)
    assert torch.equal(gslot, gslot2) and torch.equal(rows, rows2)
    assert int(t.stats[0].item()) == len(keys)
    # gather by slot
    assert torch.equal(t.gather_slots(gslot), rows)
    # read-only pull of absent keys: init values, invalid slot, nothing inserted
    absent = np.setdiff1d(_keys(500, seed=4), keys)
    rows3, gslot3 = t.pull(_dev(absent, device), insert=False)
    assert np.all(to_u32_numpy(gslot3) == 0xFFFFFFFF)
    assert np.array_equal(rows3.cpu().numpy(), init_rows(absent, dim, 99, 0.5))
    assert t.size() == len(keys)


@pytest.mark.parametrize("num_shards", [1, 4])
def test_pull_slot_cache_matches_uncached(device, num_shards):
    """ssb_pull_cached: with a worker-side key->slot cache the pull returns exactly what the probing pull
    returns — on the filling pass, on an all-hit pass, with a tiny cache that thrashes, after pushes, and
    for keys inserted later."""
    from swiftsnails_b200.ops._util import to_u32_numpy
    from swiftsnails_b200.ops.table import ShardedTable

    dim = 128
    plain = ShardedTable.local(16384, dim, device, num_shards=num_shards, frag_num=64, seed=7, init_scale=0.5)
    keys = np.unique(_keys(5000, seed=21))
    kd = _dev(keys, device)
    rows0, gslot0 = plain.pull(kd)
    for expected, dense in ((len(keys), False), (64, False), (64, True)):   # roomy / thrashing (1024 entries) / key & mask index
        t = ShardedTable.local(16384, dim, device, num_shards=num_shards, frag_num=64, seed=7, init_scale=0.5)
        cache = t.enable_slot_cache(expected, dense=dense)
        assert cache.numel() == 1 << t.slot_cache_bits and bool((cache == -1).all())
        for rep in range(3):                               # fill, then hit
            rows, gslot = t.pull(kd)
            assert torch.equal(rows, rows0)
            # slot numbering depends on insertion races, so compare through the keys: every slot holds its key
            g = to_u32_numpy(gslot)
            assert len(np.unique(g)) == len(keys)
            assert np.array_equal(g // t.capacity, t.hashfrag.to_shard(keys))
        filled = int((cache != -1).sum().item())
        assert filled > 0 and filled <= len(keys)
        if expected == len(keys) and not dense:
            assert filled > 0.7 * len(keys)                 # direct-mapped: a few keys share an index
        ck = (cache[cache != -1] >> 32).cpu().numpy().astype(np.uint32)
        assert np.isin(ck, keys).all()
        # cached slots are the table's slots: gather by the cached pull's slots == rows
        assert torch.equal(t.gather_slots(gslot), rows0)
        # updates are seen through the cache (it caches slots, not values)
        grads = torch.randn(len(keys), dim, device=device)
        t.push_sgd(gslot, grads.clone(), lr=0.5, zero_grads=False)
        rows1, gslot1 = t.pull(kd)
        assert torch.equal(gslot1, gslot)
        assert torch.allclose(rows1, rows0 - 0.5 * grads, atol=1e-6)
        # keys that arrive later are inserted and then cached; absent keys with insert=False are not cached
        more = np.setdiff1d(np.unique(_keys(1500, seed=22)), keys)
        absent = np.setdiff1d(np.unique(_keys(500, seed=23)), np.concatenate([keys, more]))
        _, ga = t.pull(_dev(absent, device), insert=False)
        assert np.all(to_u32_numpy(ga) == 0xFFFFFFFF)
        r_a, g_a = t.pull(_dev(more, device))
        r_b, g_b = t.pull(_dev(more, device))
        assert torch.equal(r_a, r_b) and torch.equal(g_a, g_b)
        assert t.size() == len(keys) + len(more)
        _, ga2 = t.pull(_dev(absent, device), insert=False)
        assert np.all(to_u32_numpy(ga2) == 0xFFFFFFFF)
        # resolve() shares the cache
        assert torch.equal(t.resolve(kd), gslot)


def test_pull_slot_cache_dense_ids(device):
    """Dense word ids with the identity index: every key gets its own entry, second pull is all hits."""
    from swiftsnails_b200.ops.table import ShardedTable

    t = ShardedTable.local(16384, 128, device, num_shards=2, frag_num=64, seed=3, init_scale=0.5, with_shadow=True)
    ids = np.random.default_rng(1).permutation(6000).astype(np.uint32)
    cache = t.enable_slot_cache(6000, dense=True)
    rows_a, gslot_a = t.pull(_dev(ids, device))
    assert int((cache != -1).sum().item()) == 6000
    inserts = int(t.stats[0].item())
    out16 = torch.empty(6000, 128, dtype=torch.bfloat16, device=device)
    gz = torch.ones(6000, 128, device=device)
    rows_b, gslot_b = t.pull(_dev(ids, device), rows_bf16=out16, grads_zero=gz)      # fp32 + bf16 + grad reset, all from cache hits
    assert torch.equal(rows_a, rows_b) and torch.equal(gslot_a, gslot_b) and int(t.stats[0].item()) == inserts
    assert torch.equal(out16, rows_a.to(torch.bfloat16)) and float(gz.abs().max()) == 0.0
    out16b = torch.zeros_like(out16)
    t.pull(_dev(ids, device), rows_bf16=out16b, want_fp32=False)                      # bf16-only: served from the bf16 mirror
    assert torch.equal(out16b, out16)
    # odd sizes: n not a multiple of 32 or of the group size
    for n in (1, 3, 31, 33, 4097):
        r, g = t.pull(_dev(ids[:n], device))
        assert torch.equal(r, rows_a[:n]) and torch.equal(g, gslot_a[:n])


def test_pull_duplicate_keys_same_batch(device):
    """Concurrent warps inserting the same key must agree on one slot (claim protocol)."""
    from swiftsnails_b200.ops._util import to_u32_numpy
    from swiftsnails_b200.ops.table import ShardedTable

    t = ShardedTable.local(8192, 32, device, num_shards=2, frag_num=16)
    keys = _keys(20_000, hi=1500, seed=5)
    rows, gslot = t.pull(_dev(keys, device))
    g = to_u32_numpy(gslot)
    for k in np.unique(keys)[:200]:
        assert len(np.unique(g[keys == k])) == 1
    assert t.size() == len(np.unique(keys))
    r = rows.cpu().numpy()
    first = {}
    for i, k in enumerate(keys[:5000]):
        if k in first:
            assert np.array_equal(r[i], r[first[k]])
        first.setdefault(k, i)


@pytest.mark.parametrize("dim", [128, 64, 1])
def test_push_sgd(device, dim):
    from swiftsnails_b200.ops.table import ShardedTable

    t = ShardedTable.local(4096, dim, device, num_shards=2, frag_num=8)
    keys = np.unique(_keys(2000, seed=6))
    kd = _dev(keys, device)
    rows, gslot = t.pull(kd)
    grads = torch.randn(len(keys), dim, device=device)
    g0 = grads.clone()
    t.push_sgd(gslot, grads, lr=0.1, zero_grads=True)
    rows2, _ = t.pull(kd)
    assert torch.allclose(rows2, rows - 0.1 * g0, rtol=1e-6, atol=1e-7)
    assert float(grads.abs().max()) == 0.0


@pytest.mark.parametrize("dim", [128, 1])
def test_push_adagrad(device, dim):
    from swiftsnails_b200.ops.table import ShardedTable
    from swiftsnails_b200.utils.reference import adagrad_ref

    t = ShardedTable.local(4096, dim, device, num_shards=2, frag_num=8, with_state=True)
    keys = np.unique(_keys(2000, seed=7))
    kd = _dev(keys, device)
    w, gslot = t.pull(kd)
    acc = torch.zeros_like(w)
    for it in range(3):
        g = torch.randn(len(keys), dim, device=device)
        t.push_adagrad(gslot, g.clone(), lr=0.05, eps=1e-8, zero_grads=False)
        w, acc = adagrad_ref(w, acc, g, 0.05, 1e-8)
    w2, _ = t.pull(kd)
    assert torch.allclose(w2, w, rtol=1e-4, atol=1e-6)


def test_neg_sample_distribution(device):
    from swiftsnails_b200.ops._util import to_u32_numpy
    from swiftsnails_b200.ops.sampling import AliasSampler

    w = 1.0 / np.arange(1, 501) ** 0.75
    s = AliasSampler(w, device, seed=5)
    out = to_u32_numpy(s.sample(2_000_001, key_offset=1000))
    assert out.min() >= 1000 and out.max() < 1500
    freq = np.bincount(out - 1000, minlength=500) / out.size
    p = w / w.sum()
    assert np.abs(freq - p).max() < 2e-3
    out2 = to_u32_numpy(s.sample(1000))
    assert not np.array_equal(out2, out[:1000] - 1000)  # counter advanced


def test_make_pairs(device):
    from swiftsnails_b200.ops import sgns
    from swiftsnails_b200.ops._util import to_u32_numpy
    from swiftsnails_b200.utils.reference import make_pairs_ref

    n, sent_len, window = 5000, 37, 5
    inv_in = np.arange(n, dtype=np.uint32)
    inv_out = np.arange(n, dtype=np.uint32) + 100_000
    inv_in[::97] = 0xFFFFFFFF
    inv_out[::89] = 0xFFFFFFFF
    pi = torch.empty(n * 2 * window, dtype=torch.int32, device=device)
    po = torch.empty_like(pi)
    cnt = torch.zeros(1, dtype=torch.int32, device=device)
    sgns.make_pairs(_dev(inv_in, device), _dev(inv_out, device), sent_len, window, 42, 1000, pi, po, cnt)
    m = int(cnt.item())
    got = sorted(zip(to_u32_numpy(pi)[:m].tolist(), to_u32_numpy(po)[:m].tolist()))
    ref = sorted(make_pairs_ref(inv_in, inv_out, sent_len, window, 42, 1000))
    assert got == ref


@pytest.mark.parametrize("dim,n_neg,n_pairs", [(128, 16, 1000), (64, 8, 300), (128, 4, 128), (256, 16, 200)])
def test_sgns_tiles_vs_torch(device, dim, n_neg, n_pairs):
    from swiftsnails_b200.ops import sgns
    from swiftsnails_b200.utils.reference import sgns_tiles_ref

    torch.manual_seed(0)
    U = 500
    params = (torch.rand(U, dim, device=device) - 0.5) * 0.8
    grads = torch.zeros_like(params)
    pin = torch.randint(0, U, (n_pairs,), device=device, dtype=torch.int32)
    pout = torch.randint(0, U, (n_pairs,), device=device, dtype=torch.int32)
    tiles = (n_pairs + 127) // 128
    neg = torch.randint(0, U, (tiles * n_neg,), device=device, dtype=torch.int32)
    neg[0] = pout[0]  # exercise the negative == target mask
    cnt = torch.tensor([n_pairs], dtype=torch.int32, device=device)
    loss = torch.zeros(2, device=device)
    sgns.sgns_tiles(params, grads, pin, pout, cnt, n_pairs + 500, neg, n_neg, 5.0 / n_neg, loss)
    gref, lref = sgns_tiles_ref(params, pin, pout, neg, n_neg, 5.0 / n_neg)
    assert torch.allclose(grads, gref, rtol=2e-4, atol=2e-5)
    assert abs(float(loss[0]) - lref) / lref < 1e-4
    assert float(loss[1]) == n_pairs


def test_lr_fwd_bwd(device):
    from swiftsnails_b200.ops import lr
    from swiftsnails_b200.utils.reference import lr_ref

    torch.manual_seed(1)
    U, n_ex, n_feat = 3000, 2000, 39
    params = torch.randn(U, 1, device=device) * 0.1
    inv = torch.randint(0, U, (n_ex, n_feat), device=device, dtype=torch.int32)
    x = torch.rand(n_ex, n_feat, device=device)
    y = (torch.rand(n_ex, device=device) < 0.3).float()
    grads = torch.zeros(U, 1, device=device)
    loss = torch.zeros(2, device=device)
    lr.lr_fwd_bwd(params, grads, inv, x, y, n_feat, loss)
    gref, lref = lr_ref(params, inv, x, y)
    assert torch.allclose(grads.reshape(-1), gref, rtol=1e-4, atol=1e-5)
    assert abs(float(loss[0]) - lref) / lref < 1e-4


def test_dump_and_load_roundtrip(device):
    from swiftsnails_b200.ops.table import ShardedTable

    t = ShardedTable.local(2048, 16, device, num_shards=2, frag_num=8, with_state=False)
    keys = np.unique(_keys(1500, seed=8))
    kd = _dev(keys, device)
    rows, gslot = t.pull(kd)
    t.push_sgd(gslot, torch.randn_like(rows), lr=1.0)
    rows, _ = t.pull(kd)
    dumped = {}
    for sh in t.owned:
        k, r = t.dump_shard(sh)
        for kk, rr in zip(k, r):
            dumped[int(kk)] = rr
    assert len(dumped) == len(keys)
    rc = rows.cpu().numpy()
    for i, k in enumerate(keys):
        assert np.array_equal(dumped[int(k)], rc[i])
    # load into a fresh table with a different shard count
    t2 = ShardedTable.local(4096, 16, device, num_shards=3, frag_num=9)
    t2.load(kd, rows)
    rows2, _ = t2.pull(kd, insert=False)
    assert torch.equal(rows2, rows)


def test_word2vec_loss_decreases(device):
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker, synthetic_corpus
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable, round_capacity

    cfg = Word2VecConfig(vocab=2000, dim=64, window=3, negatives=5, neg_pool=8, sent_len=50,
                         tokens_per_step=4096, lr=0.05, optimizer="adagrad")
    table = ShardedTable.local(round_capacity(2 * cfg.vocab), cfg.dim, device, num_shards=2, frag_num=16,
                               with_state=True, init_scale=1.0 / cfg.dim)
    w = Word2VecWorker(cfg, table, device)
    # structured corpus: token t is followed by (t*7+1) % vocab so there is signal to learn
    rng = np.random.default_rng(0)
    base = synthetic_corpus(cfg.vocab, cfg.tokens_per_step * 30, seed=3)
    base[1::2] = (base[0::2] * 7 + 1) % cfg.vocab
    losses = []
    for s in range(30):
        tok = u32_tensor(base[s * cfg.tokens_per_step:(s + 1) * cfg.tokens_per_step], device)
        w.reset_loss()
        w.step(tok)
        losses.append(w.mean_loss())
    assert all(np.isfinite(losses))
    assert np.mean(losses[-5:]) < 0.9 * np.mean(losses[:3]), losses


def test_word2vec_tcgen05_loss_decreases(device):
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker, synthetic_corpus
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable, round_capacity

    cfg = Word2VecConfig(vocab=2000, dim=128, window=3, negatives=5, neg_pool=16, sent_len=50,
                         tokens_per_step=4096, lr=0.05, optimizer="adagrad", tile_impl="tcgen05")
    table = ShardedTable.local(round_capacity(2 * cfg.vocab), cfg.dim, device, num_shards=2, frag_num=16,
                               with_state=True, init_scale=1.0 / cfg.dim)
    w = Word2VecWorker(cfg, table, device)
    base = synthetic_corpus(cfg.vocab, cfg.tokens_per_step * 30, seed=3)
    base[1::2] = (base[0::2] * 7 + 1) % cfg.vocab
    losses = []
    for s in range(30):
        tok = u32_tensor(base[s * cfg.tokens_per_step:(s + 1) * cfg.tokens_per_step], device)
        w.reset_loss()
        w.step(tok)
        losses.append(w.mean_loss())
    assert all(np.isfinite(losses))
    assert np.mean(losses[-5:]) < 0.9 * np.mean(losses[:3]), losses


def test_lr_worker_learns(device):
    from swiftsnails_b200.models.logistic_regression import LRConfig, LRWorker, criteo_like_batch
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable

    cfg = LRConfig(feature_bits=20, batch_size=4096, lr=0.1)
    table = ShardedTable.local(1 << 19, 1, device, num_shards=2, frag_num=16, with_state=True, init_scale=0.0)
    w = LRWorker(cfg, table, device)
    losses = []
    for s in range(40):
        k, v, y = criteo_like_batch(cfg, cfg.batch_size, seed=s)
        w.reset_loss()
        w.step(u32_tensor(k, device), torch.from_numpy(v).to(device), torch.from_numpy(y).to(device))
        losses.append(w.mean_loss())
    assert abs(losses[0] - np.log(2)) < 1e-3          # zero-initialised weights
    assert np.mean(losses[-5:]) < 0.97 * losses[0], losses


def test_nccl_baseline_single_rank_matches_p2p(device):
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable
    from swiftsnails_b200.parallel.nccl_baseline import NcclShardedTable

    t = ShardedTable.local(8192, 32, device, num_shards=1, with_state=True)
    nt = NcclShardedTable(t)
    keys = np.unique(_keys(3000, seed=21))
    kd = u32_tensor(keys, device)
    rows_p2p, gslot = t.pull(kd)
    rows_nccl, _ = nt.pull(kd)
    assert torch.equal(rows_p2p, rows_nccl)
    g = torch.randn(len(keys), 32, device=device)
    nt.push_sgd(None, g.clone(), lr=0.1)
    rows2, _ = t.pull(kd)
    assert torch.allclose(rows2, rows_p2p - 0.1 * g, rtol=1e-6, atol=1e-7)


def test_word2vec_pipelined_two_streams(device):
    """dedup+pull one step ahead of SGNS+push on a second stream: still learns, nothing races on
    the double-buffered caches, and flush() makes the caller's stream see the result."""
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker, synthetic_corpus
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable, round_capacity

    cfg = Word2VecConfig(vocab=2000, dim=128, window=3, negatives=5, neg_pool=16, sent_len=50,
                         tokens_per_step=4096, lr=0.05, optimizer="adagrad", tile_impl="tcgen05")
    table = ShardedTable.local(round_capacity(2 * cfg.vocab), cfg.dim, device, num_shards=2, frag_num=16,
                               with_state=True, init_scale=1.0 / cfg.dim)
    w = Word2VecWorker(cfg, table, device, pipeline=True)
    base = synthetic_corpus(cfg.vocab, cfg.tokens_per_step * 40, seed=3)
    base[1::2] = (base[0::2] * 7 + 1) % cfg.vocab
    toks = [u32_tensor(base[s * cfg.tokens_per_step:(s + 1) * cfg.tokens_per_step], device) for s in range(40)]
    losses = []
    for blk in range(8):
        w.reset_loss()
        for s in range(5):
            w.step(toks[blk * 5 + s])
        losses.append(w.mean_loss())
    assert all(np.isfinite(losses))
    assert losses[-1] < 0.9 * losses[0], losses
    assert float(w.loss[1]) > 0


def test_bf16_row_mirror_stays_in_step(device):
    """With a bf16 mirror in the shard, a bf16-only pull reads the mirror; every owner-side writer
    (lazy insert, SGD / AdaGrad push, bulk load) must keep it equal to bf16(row)."""
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable

    t = ShardedTable.local(8192, 128, device, num_shards=2, frag_num=16, with_state=True, with_shadow=True)
    assert t.has_shadow
    keys = np.unique(_keys(4000, seed=31))
    kd = u32_tensor(keys, device)
    n = len(keys)

    def check():
        fp32, gslot = t.pull(kd)                                  # reads the fp32 rows
        bf = torch.empty(n, 128, dtype=torch.bfloat16, device=device)
        t.pull(kd, rows_bf16=bf, want_fp32=False)                 # reads the mirror
        assert torch.equal(bf, fp32.to(torch.bfloat16))
        return gslot

    bf0 = torch.empty(n, 128, dtype=torch.bfloat16, device=device)
    _, gslot = t.pull(kd, rows_bf16=bf0, want_fp32=False)         # insert path writes row + mirror
    check()
    t.push_sgd(gslot, torch.randn(n, 128, device=device), lr=0.1)
    check()
    t.push_adagrad(gslot, torch.randn(n, 128, device=device), lr=0.1)
    check()
    more = np.setdiff1d(_keys(500, seed=32), keys)
    t.load(u32_tensor(more, device), torch.randn(len(more), 128, device=device))
    fp32, _ = t.pull(u32_tensor(more, device), insert=False)
    bf = torch.empty(len(more), 128, dtype=torch.bfloat16, device=device)
    t.pull(u32_tensor(more, device), rows_bf16=bf, want_fp32=False, insert=False)
    assert torch.equal(bf, fp32.to(torch.bfloat16))


@pytest.mark.parametrize("adagrad", [0, 1])
def test_bf16_mirror_under_concurrent_inbox_segments(device, adagrad):
    """Hot words reach an owner from every sender in one inbox apply: 7 segments, run side by side by one
    ssb_push_bf16_seg launch, all update the SAME 64 rows. After every launch the mirror must equal bf16(row) exactly
    (the apply itself does not touch the mirror; a trailing pass over the touched slots sets it from the final rows)."""
    from swiftsnails_b200 import _lib
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable

    dim, segs, hot, seg_rows = 128, 7, 64, 64
    t = ShardedTable.local(4096, dim, device, num_shards=1, with_state=True, with_shadow=True)
    keys = np.unique(_keys(400, seed=77))[:hot]
    kd = u32_tensor(keys, device)
    _, gslot = t.pull(kd)
    slots = (gslot % t.capacity).to(torch.int32)
    all_slots = slots.repeat(segs).contiguous()                      # segment y = the same 64 slots
    counts = torch.full((segs,), seg_rows, dtype=torch.int32, device=device)
    torch.manual_seed(5)
    for it in range(60):
        g = (torch.randn(segs * seg_rows, dim, device=device) * 0.3).to(torch.bfloat16)
        _lib.call("ssb_push_bf16_seg", t._view_ref(), all_slots.data_ptr(), g.data_ptr(), seg_rows, segs, counts.data_ptr(),
                  0.05, 1e-8, adagrad, _lib.stream_ptr())
        fp32, _ = t.pull(kd, insert=False)
        bf = torch.empty(hot, dim, dtype=torch.bfloat16, device=device)
        t.pull(kd, rows_bf16=bf, want_fp32=False, insert=False)
        assert torch.equal(bf, fp32.to(torch.bfloat16)), "mirror out of step after launch %d" % it


def test_pull_bulk_copy_variant_passes_the_pull_tests(device):
    """SSB200_PULL_BULK=1 routes pulls through pull_bulk_kernel (cp.async.bulk rows into shared memory,
    lane-per-key resolve). The switch is read once per process, so the pull / slot-cache / push tests are
    re-run in a child process with it set."""
    import os
    import subprocess
    import sys

    env = dict(os.environ, SSB200_PULL_BULK="1")
    here = os.path.abspath(__file__)
    p = subprocess.run([sys.executable, "-m", "pytest", here, "-q", "-x", "-m", "gpu", "-p", "no:cacheprovider",
                        "-k", "(pull or slot or push) and not bulk_copy_variant"],
                       capture_output=True, text=True, timeout=900, env=env, cwd=os.path.dirname(os.path.dirname(here)))
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
    assert " passed" in p.stdout
