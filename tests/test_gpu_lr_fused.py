"""The fused sparse-LR step (csrc/kernels/lr.cu, K6 fused: resolve -> gather once per touched key -> forward/backward
-> update rule once per touched key, one cooperative kernel) against a plain fp64 oracle of minibatch logistic
regression with the same update rule, and against the unfused four-launch path.
ref: core/framework/SwiftWorker.h:19-57 (loop shape), parameter/sparse_access_method.h:33-46 (server-side rule)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _batch(rng, n_ex, n_feat, space, n_numeric=3, empty_frac=0.05):
    keys = (space * rng.random((n_ex, n_feat)) ** 3).astype(np.int64) + 16          # skewed: hot and cold features
    keys[:, :n_numeric] = np.arange(n_numeric)[None, :]                           # fields every example shares
    x = np.ones((n_ex, n_feat), dtype=np.float32)
    x[:, :n_numeric] = np.log1p(rng.pareto(1.5, (n_ex, n_numeric))).astype(np.float32)
    x[rng.random((n_ex, n_feat)) < 0.02] = 0.0                                      # zero-valued features
    keys = keys.astype(np.uint32)
    keys[rng.random((n_ex, n_feat)) < empty_frac] = 0xFFFFFFFF                     # padded positions
    y = (rng.random(n_ex) < 0.4).astype(np.float32)
    return keys, x, y


def _oracle_step(w, st, keys, x, y, rule, hy):
    """w: dict key -> float64 weight; st: dict key -> state tuple."""
    from swiftsnails_b200.utils.reference import ftrl_ref

    valid = keys != 0xFFFFFFFF
    wv = np.vectorize(lambda k: w.get(int(k), 0.0))(keys) * valid
    z = (wv * x).sum(1)
    p = 1.0 / (1.0 + np.exp(-z))
    loss = float(np.sum(np.logaddexp(0, -z) * y + np.logaddexp(0, z) * (1 - y)))
    g = (p - y)[:, None] * x * valid
    uk, inv = np.unique(keys[valid], return_inverse=True)
    gs = np.zeros(len(uk))
    np.add.at(gs, inv, g[valid])
    for k, gk in zip(uk, gs):
        k = int(k)
        wk = w.get(k, 0.0)
        if gk == 0.0:
            w.setdefault(k, wk)
            continue
        if rule == "sgd":
            w[k] = wk - hy["lr"] * gk
        elif rule == "adagrad":
            a = st.get(k, 0.0) + gk * gk
            st[k] = a
            w[k] = wk - hy["lr"] * gk / np.sqrt(a + hy["eps"])
        else:
            zz, nn = st.get(k, (0.0, 0.0))
            wn, zn, n2 = ftrl_ref(np.float64(wk), np.float64(zz), np.float64(nn), np.float64(gk), hy["lr"], hy["eps"], hy["a"], hy["b"])
            w[k], st[k] = float(wn), (float(zn), float(n2))
    return loss


@pytest.mark.parametrize("rule,hy,words", [("sgd", dict(lr=0.01), 0), ("adagrad", dict(lr=0.05, eps=1e-6), 1),
                                           ("ftrl", dict(lr=0.1, eps=1.0, a=0.05, b=0.01), 2)])
@pytest.mark.parametrize("use_map,feature_major", [(True, False), (False, False), (True, True)])
def test_fused_step_matches_the_minibatch_oracle(device, rule, hy, words, use_map, feature_major):
    from swiftsnails_b200.ops import lr as lr_ops
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable

    n_ex, n_feat, bits = 3000 + 7, 11, 16                      # a tail chunk of 7 examples
    t = ShardedTable.local(1 << 16, 1, device, num_shards=3, frag_num=12, with_state=words, init_kind=1)
    state = lr_ops.FusedLRState(t, n_ex, n_feat, bits if use_map else 0, device)
    loss = torch.zeros(2, device=device)
    rng = np.random.default_rng(0)
    w, st = {}, {}
    seen = set()
    for step in range(4):
        keys, x, y = _batch(rng, n_ex, n_feat, (1 << bits) - 32)
        want_loss = _oracle_step(w, st, keys, x, y, rule, hy)
        loss.zero_()
        kk, xx = (np.ascontiguousarray(keys.T), np.ascontiguousarray(x.T)) if feature_major else (keys, x)
        lr_ops.lr_fused_step(state, u32_tensor(kk.reshape(-1), device), torch.from_numpy(xx).to(device),
                             torch.from_numpy(y).to(device), loss, rule=rule, feature_major=feature_major, **hy)
        torch.cuda.synchronize()
        c = state.counters.tolist()
        uniq = np.unique(keys[keys != 0xFFFFFFFF])
        seen |= set(int(k) for k in uniq)
        tl = state.touched[:c[0]]
        assert int((tl != -1).sum().item()) == len(uniq) and c[2] == 0 and c[3] == 0, c      # holes = lost CAS races
        got_loss, n = loss.tolist()
        assert n == n_ex and abs(got_loss - want_loss) < 2e-4 * want_loss
    all_keys = np.array(sorted(seen), dtype=np.uint32)
    assert t.size() == len(all_keys)
    got, _ = t.pull(u32_tensor(all_keys, device), insert=False)
    want = np.array([w[int(k)] for k in all_keys])
    assert np.allclose(got.cpu().numpy().reshape(-1), want, rtol=2e-3, atol=2e-5)
    assert float(state.gc.abs().max()) == 0.0                                          # gradient sums are back to zero


def test_fused_and_unfused_workers_train_alike(device):
    from swiftsnails_b200.models.logistic_regression import LRConfig, LRWorker, criteo_like_batch
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.parallel.cluster import make_table

    curves = []
    for fused in (False, True):
        cfg = LRConfig(feature_bits=20, batch_size=8192, fused=fused, lr=0.05)
        table = make_table(1 << 18, 1, device, with_state=True, init_scale=0.0)
        wk = LRWorker(cfg, table, device)
        curve = []
        for s in range(12):
            k, v, y = criteo_like_batch(cfg, cfg.batch_size, seed=s % 4)
            wk.reset_loss()
            wk.step(u32_tensor(k, device), torch.from_numpy(v).to(device), torch.from_numpy(y).to(device))
            curve.append(wk.mean_loss())
        h = wk.health()
        assert h["table_full"] == 0 and h.get("touched_overflow", 0) == 0
        curves.append(curve)
    assert curves[0][-1] < 0.97 * curves[0][0]
    assert np.allclose(curves[0], curves[1], rtol=2e-4), curves
