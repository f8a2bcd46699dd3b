"""Plain NumPy / PyTorch fp32 restatements of the device ops — the numerics oracles the GPU
tests compare against ("GPU tests comparing every kernel against a CPU oracle").
Nothing here is on a hot path."""
from __future__ import annotations

import numpy as np
import torch

from ..parallel.hashfrag import fmix64

EMPTY = 0xFFFFFFFF


def slot_hash(keys):
    k = np.asarray(keys).astype(np.uint32, copy=True)
    with np.errstate(over="ignore"):
        h = k * np.uint32(0x9E3779B1)
        h ^= h >> np.uint32(16)
        h *= np.uint32(0x85EBCA6B)
        h ^= h >> np.uint32(13)
        h *= np.uint32(0xC2B2AE35)
        h ^= h >> np.uint32(16)
    return h


def init_rows(keys, dim, seed, scale):
    """Deterministic lazy-init rows, bit-exact with ``ssb::init_value`` (common.cuh)."""
    keys = np.asarray(keys).astype(np.uint64)
    j = np.arange(dim, dtype=np.uint64)[None, :]
    x = ((keys[:, None] << np.uint64(32)) | j) ^ np.uint64(seed)
    h = fmix64(x)
    u = (h >> np.uint64(40)).astype(np.uint32).astype(np.float32) * np.float32(1.0 / 16777216.0)
    return ((u - np.float32(0.5)) * np.float32(scale)).astype(np.float32)


def philox4x32(seed, ctr):
    """Philox4x32-10 matching ``ssb::Philox::gen``. ctr: uint64 array. Returns uint32 [n,4]."""
    ctr = np.asarray(ctr, dtype=np.uint64)
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    W0, W1 = 0x9E3779B9, 0xBB67AE85
    mask = np.uint64(0xFFFFFFFF)
    c0 = ctr & mask
    c1 = ctr >> np.uint64(32)
    c2 = np.full_like(ctr, 0x5353425F)
    c3 = np.full_like(ctr, 0x42323030)
    k0 = seed & 0xFFFFFFFF
    k1 = (seed >> 32) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & mask
        hi1, lo1 = p1 >> np.uint64(32), p1 & mask
        n0 = hi1 ^ c1 ^ np.uint64(k0)
        n2 = hi0 ^ c3 ^ np.uint64(k1)
        c0, c1, c2, c3 = n0, lo1, n2, lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return np.stack([c0, c1, c2, c3], axis=1).astype(np.uint32)


def make_pairs_ref(inv_in, inv_out, sent_len, window, seed, offset):
    """Reference skip-gram pair generation (dynamic window). Returns sorted list of (in,out)."""
    n = len(inv_in)
    r = philox4x32(seed, np.arange(n, dtype=np.uint64) + np.uint64(offset))[:, 0]
    pairs = []
    for i in range(n):
        if inv_in[i] == EMPTY:
            continue
        eff = window - int(r[i] % window)
        pos = i % sent_len
        lo = -min(eff, pos)
        hi = min(eff, sent_len - 1 - pos)
        if i + hi >= n:
            hi = n - 1 - i
        for j in range(lo, hi + 1):
            if j != 0 and inv_out[i + j] != EMPTY:
                pairs.append((int(inv_in[i]), int(inv_out[i + j])))
    return pairs


def sgns_tiles_ref(params, pairs_in, pairs_out, neg_idx, n_neg, neg_weight, tile=128):
    """fp32 PyTorch reference of the SGNS tile op. Returns (grads, loss_sum)."""
    params = params.detach().to(torch.float32)
    grads = torch.zeros_like(params)
    n_pairs = pairs_in.numel()
    loss = torch.zeros((), dtype=torch.float64, device=params.device)
    for t in range((n_pairs + tile - 1) // tile):
        sl = slice(t * tile, min((t + 1) * tile, n_pairs))
        ri, ro = pairs_in[sl].long(), pairs_out[sl].long()
        nr = neg_idx[t * n_neg:(t + 1) * n_neg].long()
        a, p, ng = params[ri], params[ro], params[nr]
        s_pos = (a * p).sum(-1)
        s_neg = a @ ng.t()
        mask = (nr[None, :] != ro[:, None]).to(torch.float32)
        g_pos = torch.sigmoid(s_pos) - 1.0
        g_neg = neg_weight * torch.sigmoid(s_neg) * mask
        loss += torch.nn.functional.softplus(-s_pos).double().sum()
        loss += (neg_weight * torch.nn.functional.softplus(s_neg) * mask).double().sum()
        grads.index_add_(0, ri, g_pos[:, None] * p + g_neg @ ng)
        grads.index_add_(0, ro, g_pos[:, None] * a)
        grads.index_add_(0, nr, g_neg.t() @ a)
    return grads, float(loss)


def lr_ref(params, inv, x, labels):
    """fp32 PyTorch reference of logistic regression fwd/bwd on a dim-1 cache."""
    w = params.reshape(-1)[inv.long()]
    z = (w * x).sum(-1)
    p = torch.sigmoid(z)
    loss = torch.nn.functional.binary_cross_entropy_with_logits(z, labels, reduction="sum")
    g = (p - labels)[:, None] * x
    grads = torch.zeros_like(params.reshape(-1))
    grads.index_add_(0, inv.reshape(-1).long(), g.reshape(-1))
    return grads, float(loss)


def adagrad_ref(w, acc, g, lr, eps):
    acc = acc + g * g
    w = w - lr * g / torch.sqrt(acc + eps)
    return w, acc


def ftrl_ref(w, z, n, g, alpha, beta, l1, l2):
    """One FTRL-proximal step (McMahan et al. 2013) on float64 numpy arrays; returns (w, z, n)."""
    import numpy as np

    n_new = n + g * g
    sigma = (np.sqrt(n_new) - np.sqrt(n)) / alpha
    z_new = z + g - sigma * w
    w_new = np.where(np.abs(z_new) <= l1, 0.0, -(z_new - np.sign(z_new) * l1) / ((beta + np.sqrt(n_new)) / alpha + l2))
    return w_new, z_new, n_new


def sgns_block_ref(params, inv_in, inv_out, neg_idx, sent_len, window, seed, offset, neg_k, tile=128, halo=8,
                   n_neg=16):
    """fp32 PyTorch reference of the block-formulation SGNS step (csrc/kernels/sgns_tc.cu):
    every token is a centre, contexts are the tokens within its (shrunk) window in the same
    sentence, each tile of 128 centres shares 16 negatives weighted m_i*K/16.
    inv_* / neg_idx: int64 numpy arrays of cache rows (EMPTY = invalid).  Returns
    (grads, loss_sum, n_pairs)."""
    dev = params.device
    params = params.detach().to(torch.float32)
    n_tok = len(inv_in)
    grads = torch.zeros_like(params)
    r0 = philox4x32(seed, np.arange(n_tok, dtype=np.uint64) + np.uint64(offset))[:, 0]
    eff_all = window - (r0 % np.uint32(window)).astype(np.int64)
    loss = 0.0
    n_pairs = 0
    for t in range((n_tok + tile - 1) // tile):
        t0 = t * tile
        rows = np.arange(t0, min(t0 + tile, n_tok))
        ctx_tok = np.arange(t0 - halo, t0 + tile + halo)
        in_rows = inv_in[rows]
        ctx_ok = (ctx_tok >= 0) & (ctx_tok < n_tok)
        ctx_rows = np.where(ctx_ok, inv_out[np.clip(ctx_tok, 0, n_tok - 1)], EMPTY)
        ctx_ok &= ctx_rows != EMPTY
        rel = ctx_tok[None, :] - rows[:, None]
        mask = (rel != 0) & (np.abs(rel) <= eff_all[rows][:, None]) & ctx_ok[None, :] & \
               ((ctx_tok[None, :] // sent_len) == (rows[:, None] // sent_len)) & (in_rows != EMPTY)[:, None]
        negs = neg_idx[t * n_neg:(t + 1) * n_neg]
        neg_ok = negs != EMPTY
        mt = torch.from_numpy(mask).to(dev).float()
        a = params[torch.from_numpy(np.where(in_rows == EMPTY, 0, in_rows)).to(dev).long()]
        c = params[torch.from_numpy(np.where(ctx_rows == EMPTY, 0, ctx_rows)).to(dev).long()]
        ng = params[torch.from_numpy(np.where(neg_ok, negs, 0)).to(dev).long()]
        s_pos = a @ c.t()
        s_neg = a @ ng.t()
        m_i = mt.sum(1)
        w = (m_i * neg_k / n_neg)[:, None] * torch.from_numpy(neg_ok).to(dev).float()[None, :]
        g_pos = (torch.sigmoid(s_pos) - 1.0) * mt
        g_neg = w * torch.sigmoid(s_neg)
        loss += float((torch.nn.functional.softplus(-s_pos) * mt).double().sum())
        loss += float((w * torch.nn.functional.softplus(s_neg)).double().sum())
        n_pairs += int(mask.sum())
        d_in = g_pos @ c + g_neg @ ng
        d_ctx = g_pos.t() @ a
        d_neg = g_neg.t() @ a
        ok_in = torch.from_numpy(in_rows != EMPTY).to(dev)
        grads.index_add_(0, torch.from_numpy(np.where(in_rows == EMPTY, 0, in_rows)).to(dev).long()[ok_in], d_in[ok_in])
        ok_c = torch.from_numpy(ctx_ok).to(dev)
        grads.index_add_(0, torch.from_numpy(np.where(ctx_rows == EMPTY, 0, ctx_rows)).to(dev).long()[ok_c], d_ctx[ok_c])
        ok_n = torch.from_numpy(neg_ok).to(dev)
        grads.index_add_(0, torch.from_numpy(np.where(neg_ok, negs, 0)).to(dev).long()[ok_n], d_neg[ok_n])
    return grads, loss, n_pairs


def sgns_block_ref_batched(params, inv_in, inv_out, neg_idx, sent_len, window, seed, offset, neg_k, tile=128, halo=8, n_neg=16,
                           chunk_tiles=1024):
    """The same block-formulation SGNS step as :func:`sgns_block_ref`, all tiles at once with batched fp32 matmuls on
    ``params.device`` (the loop version takes a Python iteration per tile). ``inv_in`` / ``inv_out`` / ``neg_idx``: int64
    torch tensors on the device (EMPTY = invalid). Returns (grads, loss_sum, n_pairs). Used by
    benchmarks/precision_parity.py for the fp32 arm at full size; tests/test_reference_batched.py checks it against the loop."""
    dev = params.device
    params = params.detach().to(torch.float32)
    n_tok = int(inv_in.numel())
    n_tiles = (n_tok + tile - 1) // tile
    grads = torch.zeros_like(params)
    r0 = philox4x32(seed, np.arange(n_tok, dtype=np.uint64) + np.uint64(offset))[:, 0]
    eff_all = torch.from_numpy((window - (r0 % np.uint32(window)).astype(np.int64))).to(dev)
    pad = n_tiles * tile - n_tok
    inv_in_p = torch.cat([inv_in, torch.full((pad,), EMPTY, dtype=torch.int64, device=dev)]) if pad else inv_in
    eff_p = torch.cat([eff_all, torch.zeros(pad, dtype=torch.int64, device=dev)]) if pad else eff_all
    loss_sum = torch.zeros((), dtype=torch.float64, device=dev)
    n_pairs = torch.zeros((), dtype=torch.int64, device=dev)
    W = tile + 2 * halo
    for t0 in range(0, n_tiles, chunk_tiles):
        t1 = min(t0 + chunk_tiles, n_tiles)
        T = t1 - t0
        tt = torch.arange(t0, t1, device=dev)
        rows = tt[:, None] * tile + torch.arange(tile, device=dev)[None, :]                   # [T, tile] token positions
        ctx_tok = tt[:, None] * tile - halo + torch.arange(W, device=dev)[None, :]            # [T, W]
        in_rows = inv_in_p[rows]
        in_ok = in_rows != EMPTY
        ctx_in = (ctx_tok >= 0) & (ctx_tok < n_tok)
        ctx_rows = torch.where(ctx_in, inv_out[ctx_tok.clamp(0, n_tok - 1)], torch.full_like(ctx_tok, EMPTY))
        ctx_ok = ctx_in & (ctx_rows != EMPTY)
        rel = ctx_tok[:, None, :] - rows[:, :, None]                                           # [T, tile, W]
        mask = (rel != 0) & (rel.abs() <= eff_p[rows][:, :, None]) & ctx_ok[:, None, :] & \
               ((ctx_tok[:, None, :] // sent_len) == (rows[:, :, None] // sent_len)) & in_ok[:, :, None]
        mt = mask.to(torch.float32)
        negs = neg_idx[t0 * n_neg:t1 * n_neg].view(T, n_neg)
        neg_ok = negs != EMPTY
        a = params[torch.where(in_ok, in_rows, torch.zeros_like(in_rows))]                     # [T, tile, D]
        c = params[torch.where(ctx_ok, ctx_rows, torch.zeros_like(ctx_rows))]                  # [T, W, D]
        ng = params[torch.where(neg_ok, negs, torch.zeros_like(negs))]                         # [T, n_neg, D]
        s_pos = torch.bmm(a, c.transpose(1, 2))
        s_neg = torch.bmm(a, ng.transpose(1, 2))
        m_i = mt.sum(2)
        w = (m_i * (neg_k / n_neg))[:, :, None] * neg_ok[:, None, :].to(torch.float32)
        g_pos = (torch.sigmoid(s_pos) - 1.0) * mt
        g_neg = w * torch.sigmoid(s_neg)
        loss_sum += (torch.nn.functional.softplus(-s_pos) * mt).double().sum() + (w * torch.nn.functional.softplus(s_neg)).double().sum()
        n_pairs += mask.sum()
        d_in = torch.bmm(g_pos, c) + torch.bmm(g_neg, ng)
        d_ctx = torch.bmm(g_pos.transpose(1, 2), a)
        d_neg = torch.bmm(g_neg.transpose(1, 2), a)
        D = params.shape[1]
        grads.index_add_(0, in_rows[in_ok], d_in[in_ok].reshape(-1, D))
        grads.index_add_(0, ctx_rows[ctx_ok], d_ctx[ctx_ok].reshape(-1, D))
        grads.index_add_(0, negs[neg_ok], d_neg[neg_ok].reshape(-1, D))
    return grads, float(loss_sum.item()), int(n_pairs.item())
