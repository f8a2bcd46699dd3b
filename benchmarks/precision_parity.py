#!/usr/bin/env python
"""Does the bf16 tensor-core kernel train as well as fp32? Same objective, same data, same update rule, two arms:

  A  the product path: bf16 operands on tcgen05 (csrc/kernels/sgns_tc.cu), fp32 accumulation, fp32 table / AdaGrad
  B  the same block formulation in plain fp32 PyTorch (utils/reference.sgns_block_ref: fp32 matmuls, TF32 off), the
     gradients pushed with the same AdaGrad kernel

Both start from the same table (same seed -> same lazy init), see the same batches, the same window shrinks and the
same negatives (same sampler seed). After training, BOTH tables are scored on held-out batches with the fp32
reference. Output: one JSON line (loss per pair of each arm on the held-out set, relative difference, training
curves). Small shapes: the fp32 arm is a Python loop over tiles."""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def big(steps):
    """bf16 tcgen05 (product path) vs the SAME block objective in fp32 (batched PyTorch matmuls, TF32 off:
    utils/reference.sgns_block_ref_batched), same table seed, same planted-structure corpus, same windows and negatives
    (same sampler seed), same AdaGrad push kernel; >= 64 M tokens. Both tables are then scored by the SAME evaluator — the
    fp32 oracle with a DIFFERENT sampler seed (independent negatives) — on held-out batches. A third arm, the fp32
    CUDA-core pair kernel (`tile_impl="cuda_core"`, negatives shared per 128 PAIRS instead of per 128 centre tokens), shows
    what a different sampling scheme does at the same precision."""
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable, round_capacity
    from swiftsnails_b200.utils.reference import EMPTY, sgns_block_ref_batched

    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda", 0)
    V, B = 200_000, 524_288
    rng = np.random.default_rng(0)
    partner = rng.permutation(V)

    def batch():
        left = (V * rng.random(B // 2) ** 2).astype(np.int64)
        t = np.empty(B, dtype=np.uint32)
        t[0::2], t[1::2] = left, partner[left]
        return t

    n_distinct = 16
    train = [u32_tensor(batch(), dev) for _ in range(n_distinct)]       # cycled: the table keeps training
    held = [u32_tensor(batch(), dev) for _ in range(4)]

    def cfg_of(impl, lr=0.05, seed=7):
        return Word2VecConfig(vocab=V, dim=128, window=5, negatives=5, neg_pool=16, sent_len=1024, tokens_per_step=B, lr=lr,
                              optimizer="adagrad", tile_impl=impl, seed=seed)

    def ref_step(w, s, step_idx):
        """fp32 oracle gradients / loss for the batch that _comm() prepared in buffer set s (device tensors throughout)."""
        inv = s.inverse.to(torch.int64) & 0xFFFFFFFF
        nu = int(s.n_unique.item())
        return sgns_block_ref_batched(s.params[:max(nu, 1)], inv[:B], inv[B:2 * B], inv[2 * B:], w.cfg.sent_len, w.cfg.window,
                                      w.cfg.seed, step_idx * B, float(w.cfg.negatives))

    def heldout(table):
        we = Word2VecWorker(cfg_of("cuda_core", lr=0.0, seed=1234), table, dev, pipeline=False)    # pulls fp32 rows, other negatives
        tot, cnt = 0.0, 0
        for i, t in enumerate(held):
            s = we.sets[0]
            s.step_idx = 5000 + i
            we._comm(s, t)
            _, loss, n_pairs = ref_step(we, s, s.step_idx)
            tot += loss
            cnt += n_pairs
        return tot / max(cnt, 1)

    out = {"bench": "precision_parity_big", "vocab": V, "tokens_per_step": B, "steps": steps, "tokens": steps * B}
    score, curves = {}, {}
    for arm in ("bf16_tcgen05", "fp32_same_objective", "fp32_cuda_core_pair_kernel"):
        impl = {"bf16_tcgen05": "tcgen05", "fp32_same_objective": "cuda_core", "fp32_cuda_core_pair_kernel": "cuda_core"}[arm]
        table = ShardedTable.local(round_capacity(2 * V + 64), 128, dev, with_state=True, init_scale=1.0 / 128, seed=11)
        w = Word2VecWorker(cfg_of(impl), table, dev, pipeline=False)
        curve = []
        for st in range(steps):
            if arm == "fp32_same_objective":
                s = w.sets[0]
                s.step_idx = st
                w._comm(s, train[st % n_distinct])
                g, loss, n_pairs = ref_step(w, s, st)
                s.grads[:g.shape[0]].copy_(g)
                w._push(s)
                cur = loss / max(n_pairs, 1)
            else:
                w.reset_loss()
                w.step(train[st % n_distinct])
                cur = w.mean_loss() if (st % 16 == 0 or st == steps - 1) else None
            if st % 16 == 0 or st == steps - 1:
                curve.append(round(cur, 5))
        curves[arm] = curve
        score[arm] = heldout(table)
        del w, table
        torch.cuda.empty_cache()
    out["heldout_loss_per_pair_fp32_evaluator_independent_negatives"] = score
    out["relative_difference_bf16_vs_fp32_same_objective"] = abs(score["bf16_tcgen05"] - score["fp32_same_objective"]) / score["fp32_same_objective"]
    out["train_loss_curve_every_16_steps"] = curves
    out["untrained_loss_per_pair"] = float(np.log(2.0) * 6)
    print(json.dumps(out))


def main():
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker
    from swiftsnails_b200.ops._util import to_u32_numpy, u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable, round_capacity
    from swiftsnails_b200.utils.reference import EMPTY, sgns_block_ref

    ap = argparse.ArgumentParser()
    ap.add_argument("--vocab", type=int, default=4000)
    ap.add_argument("--tokens", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=120)
    ap.add_argument("--heldout", type=int, default=6)
    ap.add_argument("--big", type=int, default=0, help="the large run: bf16 tcgen05 kernel vs the fp32 CUDA-core kernel, N steps of "
                    "524 288 tokens (128 steps = 67 M tokens), held-out loss of both tables from the fp32 kernel with independent negatives")
    a = ap.parse_args()
    if a.big:
        return big(a.big)
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda", 0)
    V, B = a.vocab, a.tokens
    rng = np.random.default_rng(0)
    partner = rng.permutation(V)

    def batch():
        # word a is followed by partner(a): structure to learn, Zipf-ish unigram
        left = (V * rng.random(B // 2) ** 2).astype(np.int64)
        t = np.empty(B, dtype=np.uint32)
        t[0::2], t[1::2] = left, partner[left]
        return t

    train = [batch() for _ in range(a.steps)]
    held = [batch() for _ in range(a.heldout)]

    def make(impl):
        cfg = Word2VecConfig(vocab=V, dim=128, window=5, negatives=5, neg_pool=16, sent_len=64, tokens_per_step=B, lr=0.05,
                             optimizer="adagrad", tile_impl=impl, seed=7)
        table = ShardedTable.local(round_capacity(2 * V + 64), 128, dev, with_state=True, init_scale=1.0 / 128, seed=11)
        return cfg, table, Word2VecWorker(cfg, table, dev, pipeline=False)

    def ref_on(w, s, tokens):
        """fp32 reference gradients / loss for the batch that _comm() prepared in buffer set s."""
        cfg = w.cfg
        inv = to_u32_numpy(s.inverse).astype(np.int64)
        inv[inv == 0xFFFFFFFF] = EMPTY
        nu = int(s.n_unique.item())
        params = s.params[:max(nu, 1)]
        return sgns_block_ref(params, inv[:B], inv[B:2 * B], inv[2 * B:], cfg.sent_len, cfg.window, cfg.seed,
                              s.step_idx * B, float(cfg.negatives)), nu

    curves = {}
    tables = {}
    # arm A: the tensor-core kernel
    cfgA, tabA, wA = make("tcgen05")
    ca = []
    for t in train:
        wA.reset_loss()
        wA.step(u32_tensor(t, dev))
        ca.append(wA.mean_loss())
    curves["bf16_tcgen05"], tables["bf16_tcgen05"] = ca, tabA
    # arm B: fp32 PyTorch gradients, same push
    cfgB, tabB, wB = make("cuda_core")
    cb = []
    for i, t in enumerate(train):
        s = wB.sets[0]
        s.step_idx = i
        wB._comm(s, u32_tensor(t, dev))
        (g, loss, n_pairs), nu = ref_on(wB, s, t)
        s.grads[:g.shape[0]].copy_(g)
        wB._push(s)
        cb.append(loss / max(n_pairs, 1))
    curves["fp32_reference"], tables["fp32_reference"] = cb, tabB
    # held-out score of both tables with the fp32 reference (no update)
    score = {}
    for name, tab in tables.items():
        cfgE = Word2VecConfig(vocab=V, dim=128, window=5, negatives=5, neg_pool=16, sent_len=64, tokens_per_step=B, lr=0.0,
                              optimizer="adagrad", tile_impl="cuda_core", seed=99)
        wE = Word2VecWorker(cfgE, tab, dev, pipeline=False)
        tot, cnt = 0.0, 0
        for i, t in enumerate(held):
            s = wE.sets[0]
            s.step_idx = 1000 + i
            wE._comm(s, u32_tensor(t, dev))
            (g, loss, n_pairs), _ = ref_on(wE, s, t)
            tot += loss
            cnt += n_pairs
        score[name] = tot / max(cnt, 1)
    rel = abs(score["bf16_tcgen05"] - score["fp32_reference"]) / score["fp32_reference"]
    untrained = float(np.log(2.0) * (1 + 5))
    print(json.dumps({"bench": "precision_parity", "vocab": V, "tokens_per_step": B, "steps": a.steps,
                      "heldout_loss_per_pair": score, "relative_difference": rel, "untrained_loss_per_pair": untrained,
                      "train_loss_first_last": {k: [v[0], v[-1]] for k, v in curves.items()},
                      "train_curve_max_rel_diff": float(np.max(np.abs(np.array(ca) - np.array(cb)) / np.array(cb))),
                      "arms": {"bf16_tcgen05": "bf16 operands, tcgen05.mma, fp32 accumulate; fp32 table + AdaGrad state",
                               "fp32_reference": "same block objective in fp32 PyTorch (TF32 off), same AdaGrad push kernel"}}))


if __name__ == "__main__":
    main()
