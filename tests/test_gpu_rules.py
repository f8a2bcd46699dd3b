"""Device-side update rules (include/ssb200_rule.cuh, ops/rules.py, framework.PushMethod).

ref: /root/reference/src/core/parameter/sparse_access_method.h:21-46 — the user's PushAccessMethod / PullAccessMethod.
Here: the built-in rules by name against fp32/fp64 oracles; FTRL (two state words, rows created as zeros) — a rule
that exists only in the public header and csrc/kernels/rules.cu; the owner-side apply of bf16 inbox segments with
"ranks" emulated in one process (the cross-GPU run is tests/test_multigpu.py); checkpoint of a two-word state; and a
user's rule compiled from a source file at run time."""
import os

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

USER_RULE = r'''
#include "ssb200_rule.cuh"
// sign-SGD with a per-weight update counter: w -= lr * sign(g) / sqrt(1 + count); count += 1
struct SignDecay {
  static constexpr int kStateWords = 1;
  static constexpr bool kSkipZeroGrad = true;
  static constexpr int kInit = SSB_INIT_CONSTANT;
  static __device__ float one(float g, float c, const ssb::rule::Hyper& h) {
    return -h.lr * (g > 0.f ? 1.f : -1.f) * rsqrtf(1.f + c);
  }
  static __device__ float4 apply4(const float*, float* st, int, float4 g, const ssb::rule::Hyper& h) {
    const float4 c = ssb::atom_add_f4_sys(st, make_float4(g.x != 0.f, g.y != 0.f, g.z != 0.f, g.w != 0.f));
    return make_float4(g.x != 0.f ? one(g.x, c.x, h) : 0.f, g.y != 0.f ? one(g.y, c.y, h) : 0.f,
                       g.z != 0.f ? one(g.z, c.z, h) : 0.f, g.w != 0.f ? one(g.w, c.w, h) : 0.f);
  }
  static __device__ float apply1(const float*, float* st, int, float g, const ssb::rule::Hyper& h) {
    return one(g, ssb::atom_add_f1_sys(st, 1.f), h);
  }
};
SSB200_EXPORT_RULE(signdecay, SignDecay)
'''


def _keys(n, seed):
    return np.unique(np.random.default_rng(seed).integers(0, 1 << 30, n).astype(np.uint32))


def test_user_rule_compiles_from_source_for_sm100a(tmp_path, native_built):
    """CPU: nvcc cross-compiles the user's file against the public header only; the info entry point answers."""
    from swiftsnails_b200.ops import rules

    src = tmp_path / "signdecay.cu"
    src.write_text(USER_RULE)
    r = rules.Rule.from_source("signdecay", str(src), force=True)
    assert r.state_words == 1 and r.init_kind == rules.INIT_CONSTANT
    assert rules.Rule.get("signdecay") is r
    ftrl = rules.Rule.get("ftrl")
    assert ftrl.state_words == 2 and ftrl.init_kind == rules.INIT_ZERO
    assert rules.Rule.get("sgd").state_words == 0 and rules.Rule.get("adagrad").state_words == 1
    with pytest.raises(KeyError):
        rules.Rule.get("no_such_rule")


@pytest.mark.gpu
@pytest.mark.parametrize("dim", [128, 6])
def test_named_sgd_and_adagrad_equal_the_stock_push(device, dim):
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable

    keys = _keys(3000, 1)
    kd = u32_tensor(keys, device)
    torch.manual_seed(0)
    gs = [torch.randn(len(keys), dim, device=device) for _ in range(3)]
    a = ShardedTable.local(4096, dim, device, num_shards=2, frag_num=8, with_state=True, with_shadow=(dim == 128), seed=3)
    b = ShardedTable.local(4096, dim, device, num_shards=2, frag_num=8, with_state=True, with_shadow=(dim == 128), seed=3)
    _, sa = a.pull(kd)
    _, sb = b.pull(kd)
    a.push_sgd(sa, gs[0].clone(), lr=0.1)
    b.push_rule("sgd", sb, gs[0].clone(), lr=0.1)
    for g in gs[1:]:
        a.push_adagrad(sa, g.clone(), lr=0.05, eps=1e-6)
        b.push_rule("adagrad", sb, g.clone(), lr=0.05, eps=1e-6)
    ra, _ = a.pull(kd, insert=False)
    rb, _ = b.pull(kd, insert=False)
    assert torch.equal(ra, rb)
    if dim == 128:
        bf = torch.empty(len(keys), dim, dtype=torch.bfloat16, device=device)
        b.pull(kd, rows_bf16=bf, want_fp32=False, insert=False)
        assert torch.equal(bf, rb.to(torch.bfloat16))


@pytest.mark.gpu
@pytest.mark.parametrize("dim", [1, 64])
def test_ftrl_matches_the_float64_oracle(device, dim):
    """Sparse-LR shape (dim 1) and a vector shape; unique keys, four steps, l1 large enough to zero some weights."""
    from swiftsnails_b200.framework import PushMethod
    from swiftsnails_b200.ops import rules
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable
    from swiftsnails_b200.utils.reference import ftrl_ref

    m = PushMethod("ftrl", lr=0.1, eps=1.0, a=0.02, b=0.001)
    assert m.state_words == 2 and m.init_kind == rules.INIT_ZERO
    t = ShardedTable.local(8192, dim, device, num_shards=3, frag_num=12, with_state=m.state_words, init_kind=m.init_kind)
    keys = _keys(5000, 2)
    kd = u32_tensor(keys, device)
    rows, gslot = t.pull(kd)
    assert float(rows.abs().max()) == 0.0                     # the rule's init kind: zeros
    w = np.zeros((len(keys), dim)); z = np.zeros_like(w); n = np.zeros_like(w)
    rng = np.random.default_rng(0)
    for step in range(4):
        g = (rng.standard_normal((len(keys), dim)) * 0.05).astype(np.float32)
        g[rng.random(g.shape) < 0.3] = 0.0
        gd = torch.from_numpy(g).to(device)
        t.push_rule(m.rule, gslot, gd, zero_grads=True, **m.hyper)
        assert float(gd.abs().max()) == 0.0
        w, z, n = ftrl_ref(w, z, n, g.astype(np.float64), 0.1, 1.0, 0.02, 0.001)
    got, _ = t.pull(kd, insert=False)
    got = got.cpu().numpy()
    assert int((~np.isclose(got, w, rtol=2e-4, atol=2e-6)).sum()) <= 2
    # the l1 threshold agrees (a weight whose |z| sits within float rounding of l1 may fall on either side)
    assert (w == 0).mean() > 0.05 and int(((got == 0) != (w == 0)).sum()) <= 2
    with pytest.raises(ValueError):
        ShardedTable.local(1024, dim, device, with_state=True).push_rule("ftrl", gslot[:8], gd[:8], lr=0.1)


@pytest.mark.gpu
def test_ftrl_owner_side_apply_of_bf16_segments_two_emulated_ranks(device):
    """The multi-rank push path of a named rule, ranks emulated in one process: rank r applies its own rows with
    rule.push(only_shard=r); what it sends to the other owner arrives as a bf16 segment that the owner applies with
    rule.push_bf16_seg (+ mirror refresh). Keys are disjoint between the senders, so the oracle is exact FTRL with the
    wire rounding applied to the rows that travelled."""
    from swiftsnails_b200.ops import rules
    from swiftsnails_b200.ops._util import to_u32_numpy, u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable
    from swiftsnails_b200.parallel.hashfrag import HashFrag
    from swiftsnails_b200.utils.reference import ftrl_ref

    R, dim = 2, 128
    rule = rules.Rule.get("ftrl")
    hy = dict(lr=0.05, eps=1.0, a=0.01, b=0.0)
    t = ShardedTable.local(8192, dim, device, num_shards=R, frag_num=64, with_state=2, with_shadow=True, init_kind=rule.init_kind)
    locals_ = [ShardedTable(HashFrag(1, 1), t.capacity, dim, [t.owned[r].ptrs()], device, owned=[t.owned[r]],
                            init_kind=rule.init_kind) for r in range(R)]
    keys = _keys(6000, 3)
    parts = np.array_split(np.random.default_rng(1).permutation(keys), R)
    state = {r: [np.zeros((len(parts[r]), dim)) for _ in range(3)] for r in range(R)}      # w, z, n per sender's keys
    rng = np.random.default_rng(2)
    for step in range(3):
        seg_slots, seg_rows, counts = {}, {}, {}
        for r in range(R):                       # "send" phase of rank r
            kd = u32_tensor(parts[r], device)
            _, gslot = t.pull(kd)
            g = (rng.standard_normal((len(parts[r]), dim)) * 0.05).astype(np.float32)
            gd = torch.from_numpy(g).to(device)
            owner = to_u32_numpy(gslot) // t.capacity
            rule.push(t, gslot, gd, rule.hyper(**hy), zero_grads=False, only_shard=r)
            away = torch.from_numpy(np.nonzero(owner != r)[0]).to(device)
            seg_slots[r] = (gslot[away] % t.capacity).to(torch.int32).contiguous()
            seg_rows[r] = gd[away].to(torch.bfloat16).contiguous()
            g_eff = np.where((owner == r)[:, None], g, seg_rows_to_np(gd.to(torch.bfloat16)))
            w, z, n = state[r]
            state[r] = list(ftrl_ref(w, z, n, g_eff.astype(np.float64), 0.05, 1.0, 0.01, 0.0))
        for o in range(R):                       # "apply" phase of owner o: one segment per sender
            S = max(len(seg_slots[r]) for r in range(R))
            slots = torch.zeros(R * S, dtype=torch.int32, device=device)
            rows = torch.zeros(R * S, dim, dtype=torch.bfloat16, device=device)
            cnt = torch.zeros(R, dtype=torch.int32, device=device)
            for r in range(R):
                if r == o:
                    continue
                m = len(seg_slots[r])
                slots[r * S:r * S + m] = seg_slots[r]
                rows[r * S:r * S + m] = seg_rows[r]
                cnt[r] = m
            rule.push_bf16_seg(locals_[o], slots, rows, S, cnt, R, rule.hyper(**hy))
    for r in range(R):
        kd = u32_tensor(parts[r], device)
        got, _ = t.pull(kd, insert=False)
        assert np.allclose(got.cpu().numpy(), state[r][0], rtol=3e-4, atol=3e-6)
        bf = torch.empty(len(parts[r]), dim, dtype=torch.bfloat16, device=device)
        t.pull(kd, rows_bf16=bf, want_fp32=False, insert=False)
        assert torch.equal(bf, got.to(torch.bfloat16))                 # mirror == bf16(row) after local + inbox updates


def seg_rows_to_np(t_bf16):
    return t_bf16.to(torch.float32).cpu().numpy()


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", ["binary", "text"])
def test_checkpoint_carries_both_ftrl_state_words(device, tmp_path, fmt):
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable
    from swiftsnails_b200.utils import dump

    dim, hy = 8, dict(lr=0.1, eps=1.0, a=0.01, b=0.0)
    keys = _keys(2000, 5)
    kd = u32_tensor(keys, device)
    torch.manual_seed(1)
    g1, g2 = torch.randn(len(keys), dim, device=device) * 0.1, torch.randn(len(keys), dim, device=device) * 0.1
    a = ShardedTable.local(4096, dim, device, num_shards=2, frag_num=8, with_state=2, init_kind=1)
    _, sa = a.pull(kd)
    a.push_rule("ftrl", sa, g1.clone(), **hy)
    path = str(tmp_path / ("ck." + fmt))
    if fmt == "binary":
        dump.save_table_binary(a, path, max_bytes=32 << 10)
    else:
        dump.save_table(a, path, with_state=True, digits=9, max_bytes=32 << 10)
    a.push_rule("ftrl", sa, g2.clone(), **hy)
    want, _ = a.pull(kd, insert=False)
    b = ShardedTable.local(8192, dim, device, num_shards=3, frag_num=9, with_state=2, init_kind=1)
    assert (dump.load_table_binary if fmt == "binary" else dump.load_table)(b, path) == len(keys)
    _, sb = b.pull(kd, insert=False)
    b.push_rule("ftrl", sb, g2.clone(), **hy)
    got, _ = b.pull(kd, insert=False)
    assert torch.allclose(got, want, rtol=1e-5, atol=1e-7)


@pytest.mark.gpu
def test_user_rule_from_source_runs(device, tmp_path):
    from swiftsnails_b200.framework import DeviceParamCache, PushMethod, pull_with_barrier, push_with_barrier
    from swiftsnails_b200.ops import rules
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable

    src = tmp_path / "signdecay.cu"
    src.write_text(USER_RULE)
    rule = rules.Rule.from_source("signdecay", str(src))
    m = PushMethod(rule, lr=0.25)
    dim = 32
    t = ShardedTable.local(4096, dim, device, num_shards=2, frag_num=8, with_state=m.state_words, init_kind=m.init_kind,
                           init_scale=0.5)
    keys = _keys(1500, 6)
    cache = DeviceParamCache(4096, dim, device)
    kd = u32_tensor(keys, device)
    pull_with_barrier(t, kd, cache)
    n = int(cache.n_unique.item())
    assert n == len(keys) and torch.all(cache.params[:n] == 0.5)          # SSB_INIT_CONSTANT
    want = torch.full((n, dim), 0.5, device=device)
    torch.manual_seed(2)
    for step in range(3):
        g = torch.randn(n, dim, device=device)
        cache.grads[:n] = g
        push_with_barrier(t, cache, m)
        want -= 0.25 * torch.sign(g) / (1.0 + step) ** 0.5
        assert float(cache.grads[:n].abs().max()) == 0.0
    got, _ = t.pull(cache.unique[:n], insert=False)
    assert torch.allclose(got, want, rtol=1e-5, atol=1e-6)
