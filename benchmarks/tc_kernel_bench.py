#!/usr/bin/env python
"""Device time of the tensor-core SGNS tile kernel alone, on the real operands of a bench.py step
(Zipf corpus -> keys -> dedup -> pull into the bf16 cache), for several kernel variants:

    mode 1  pipelined kernel (default)          mode 3  same, gradient reductions skipped (experiment)
    mode 9  round-1 serial kernel               mode 11 same, reductions skipped      mode 13 serial + diagonal epilogue

Median of the timed launches (CUDA events, one launch per measurement, grads re-zeroed by the pull of
every step so each launch sees the same state). Prints one JSON line."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker, synthetic_corpus
    from swiftsnails_b200.ops import sgns as sgns_ops
    from swiftsnails_b200.parallel.cluster import make_table

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    tps = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
    modes = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 3, 9, 11, 13]
    cfg = Word2VecConfig(tile_impl="tcgen05", optimizer="adagrad", lr=0.05, tokens_per_step=tps, seed=1)
    table = make_table(2 * cfg.vocab, cfg.dim, dev, with_state=True, init_scale=1.0 / cfg.dim, frag_num=1024)
    allk = torch.arange(0, 2 * cfg.vocab, dtype=torch.int32, device=dev)
    for c in range(0, allk.numel(), 1 << 20):
        table.resolve(allk[c:c + (1 << 20)])
    w = Word2VecWorker(cfg, table, dev)
    steps = 8
    toks = torch.from_numpy(synthetic_corpus(cfg.vocab, steps * tps, seed=50).view(np.int32)).reshape(steps, -1).to(dev)
    s0, B = w.sets[0], tps
    res = {}
    for mode in modes:
        ts, loss = [], None
        for s in range(steps):
            w._comm(s0, toks[s])                     # keys, negatives, dedup, pull (zeroes the grads)
            w.loss.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            sgns_ops.sgns_block_tc(s0.params_bf16, s0.grads, s0.inverse[:B], s0.inverse[B:2 * B], cfg.sent_len, cfg.window,
                                   cfg.seed, s * B, s0.inverse[2 * B:], float(cfg.negatives), w.loss, mode=mode)
            e1.record()
            torch.cuda.synchronize()
            if s >= 2:
                ts.append(e0.elapsed_time(e1) * 1e3)
            l = w.loss.tolist()
            loss = l[0] / max(l[1], 1.0)
            gsum = float(s0.grads.abs().sum().item())
        ts.sort()
        res["mode%d" % mode] = {"us_median": round(ts[len(ts) // 2], 1), "us_min": round(ts[0], 1), "loss_per_pair": round(loss, 4),
                                "grad_abs_sum": round(gsum, 2)}
        print("mode", mode, res["mode%d" % mode], file=sys.stderr, flush=True)
    res.update({"tokens": tps, "unique_keys": int(s0.n_unique.item()), "tiles": (tps + 127) // 128})
    print(json.dumps(res))


if __name__ == "__main__":
    main()
