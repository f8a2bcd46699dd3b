#!/usr/bin/env python
"""Loss per pair as a function of tokens seen, for several tokens-per-step settings (one GPU, same
corpus, same learning rate, fresh table each time).  Progressive validation: a step's loss is
computed with the rows pulled *before* that step's update, so the mean over the last window of
tokens is an honest held-out number.  Used to pick the default minibatch of bench.py."""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecTrainer, synthetic_corpus
    from swiftsnails_b200.parallel.cluster import make_table

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    total = int(sys.argv[1]) if len(sys.argv) > 1 else 64 * 1024 * 1024
    window_tokens = 8 * 1024 * 1024
    lrs = [float(x) for x in (sys.argv[2].split(",") if len(sys.argv) > 2 else ["0.05"])]
    corpus = torch.from_numpy(synthetic_corpus(1_000_000, total, seed=11).view(np.int32)).pin_memory()
    for tps in (65536, 262144, 524288, 1048576, 2097152):
        for lr in lrs:
            cfg = Word2VecConfig(tile_impl="tcgen05", optimizer="adagrad", lr=lr, tokens_per_step=tps, seed=5)
            table = make_table(2 * cfg.vocab, cfg.dim, dev, with_state=True, init_scale=1.0 / cfg.dim)
            tr = Word2VecTrainer(cfg, table, dev, pipeline=True)
            steps = total // tps
            losses = []
            torch.cuda.synchronize()
            t0 = time.time()
            for s in range(steps):
                prev = tr.train_step(corpus[s * tps:(s + 1) * tps])
                if prev is not None:
                    losses.append(prev)
            losses.append(tr.finish())
            torch.cuda.synchronize()
            dt = time.time() - t0
            k = max(1, window_tokens // tps)
            marks = {}
            for frac in (0.25, 0.5, 1.0):
                e = int(steps * frac)
                marks["loss@%dM" % (total * frac / 2**20)] = round(float(np.mean(losses[max(0, e - k):e])), 4)
            print(json.dumps({"tokens_per_step": tps, "lr": lr, "steps": steps, **marks,
                              "wall_words_per_s": round(total / dt / 1e6, 1)}), flush=True)
            del tr, table
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
