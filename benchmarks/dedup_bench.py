#!/usr/bin/env python
"""Device time of the key de-duplication (K2) on a word2vec-shaped key batch: Zipf tokens as input
keys, the same + vocab as output keys, plus alias-sampled negatives.  Back-to-back launches between
two CUDA events."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.models.word2vec import synthetic_corpus
    from swiftsnails_b200.ops.dedup import Deduper

    dev = torch.device("cuda", 0)
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
    vocab = 1_000_000
    reps = 8
    toks = torch.from_numpy(synthetic_corpus(vocab, reps * B, seed=3).view(np.int32)).reshape(reps, B).to(dev)
    negs = torch.from_numpy(synthetic_corpus(vocab, reps * (B // 8), seed=4).view(np.int32)).reshape(reps, -1).to(dev)
    keys = torch.cat([toks, toks + vocab, negs + vocab], dim=1).contiguous()
    n = keys.shape[1]
    d = Deduper(n, dev)
    uniq = torch.empty(n, dtype=torch.int32, device=dev)
    inv = torch.empty(n, dtype=torch.int32, device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    for r in range(reps):
        d(keys[r], unique=uniq, inverse=inv, n_unique=cnt)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 10
    e0.record()
    for _ in range(iters):
        for r in range(reps):
            d(keys[r], unique=uniq, inverse=inv, n_unique=cnt)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (iters * reps)
    print(json.dumps({"bench": "dedup", "keys": n, "unique_last": int(cnt.item()), "us_per_call": round(us, 2),
                      "gkeys_per_s": round(n / us / 1e3, 2)}))


if __name__ == "__main__":
    main()
