#!/usr/bin/env python
"""Device time of the word2vec-shaped pull (unique keys of one 524 288-token Zipf batch -> bf16 cache
rows + zeroed gradient rows) on one GPU, with and without the worker-side slot cache, and of the
AdaGrad push of the same rows.  usage: pull_bench.py [slot_cache 0|1|2 (2 = dense index)] [tokens_per_step]"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.models.word2vec import synthetic_corpus
    from swiftsnails_b200.ops.dedup import Deduper
    from swiftsnails_b200.parallel.cluster import make_table

    use_cache = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 524288
    dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
    vocab, D, reps = 1_000_000, 128, 6
    table = make_table(2 * vocab, D, dev, with_state=True, init_scale=1.0 / D)
    allk = torch.arange(0, 2 * vocab, dtype=torch.int32, device=dev)
    for c in range(0, allk.numel(), 1 << 20):
        table.resolve(allk[c:c + (1 << 20)])
    if use_cache:
        table.enable_slot_cache(2 * vocab, dense=(use_cache == 2))
    toks = torch.from_numpy(synthetic_corpus(vocab, reps * B, seed=3).view(np.int32)).reshape(reps, B).to(dev)
    negs = torch.from_numpy(synthetic_corpus(vocab, reps * (B // 8), seed=4).view(np.int32)).reshape(reps, -1).to(dev)
    keys = torch.cat([toks, toks + vocab, negs + vocab], dim=1).contiguous()
    n = keys.shape[1]
    d = Deduper(n, dev)
    uniq = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(reps)]
    cnt = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(reps)]
    inv = torch.empty(n, dtype=torch.int32, device=dev)
    for r in range(reps):
        d(keys[r], unique=uniq[r], inverse=inv, n_unique=cnt[r])
    rows = torch.empty((n, D), dtype=torch.bfloat16, device=dev)
    grads = torch.empty((n, D), dtype=torch.float32, device=dev)
    gslot = torch.empty(n, dtype=torch.int32, device=dev)

    def pull(r):
        table.pull(uniq[r], n_dev=cnt[r], rows_out=None, rows_bf16=rows, want_fp32=False, gslot_out=gslot, grads_zero=grads, n=n)

    def push(r):
        table.push_adagrad(gslot, grads, 0.05, 1e-8, n_dev=cnt[r], zero_grads=False, n=n)

    for it in range(2):
        for r in range(reps):
            pull(r); push(r)
    torch.cuda.synchronize()
    res = {}
    for name, fn in (("pull", pull), ("push", push)):
        if name == "push":
            grads.normal_()            # the pull zeroed them; an all-zero gradient row is skipped by the push
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 5
        e0.record()
        for _ in range(iters):
            for r in range(reps):
                fn(r)
        e1.record(); torch.cuda.synchronize()
        res[name + "_us"] = round(e0.elapsed_time(e1) * 1e3 / (iters * reps), 1)
    nu = int(cnt[0].item())
    res.update({"bench": "pull_bench", "slot_cache": use_cache, "variant": os.environ.get("SSB200_PULL_VARIANT", "default"),
                "tokens": B, "unique_keys": nu,
                "pull_GBs": round(nu * (D * 4 + D * 2 + D * 4) / res["pull_us"] / 1e3, 1),
                "push_GBs": round(nu * (4 * D * 4 + D * 4) / res["push_us"] / 1e3, 1)})
    print(json.dumps(res))


if __name__ == "__main__":
    main()
