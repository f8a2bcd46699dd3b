#!/usr/bin/env python
"""Per-phase device time of the word2vec step (CUDA events around each op, non-pipelined), at the
launched world size: key build + negative sampling, dedup, pull, SGNS tiles, push.  Max over ranks."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker, synthetic_corpus
    from swiftsnails_b200.parallel.cluster import make_table

    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local); torch.cuda.set_device(dev)
    from swiftsnails_b200.parallel import group as ps_group
    grp = ps_group.init("master")
    rank = grp.rank
    opt = sys.argv[1] if len(sys.argv) > 1 else "adagrad"
    tps = int(sys.argv[2]) if len(sys.argv) > 2 else 524288
    cfg = Word2VecConfig(tile_impl="tcgen05", optimizer=opt, lr=0.05, tokens_per_step=tps, seed=1 + rank, sent_len=1024, slot_cache=True)
    # same table options as bench.py at every N: bf16 row mirror, worker slot cache, inbox push (bf16 rows, flag hand-off) with peers
    table = make_table((2 * cfg.vocab + world - 1) // world, cfg.dim, dev, with_state=(opt == "adagrad"), init_scale=1.0 / cfg.dim,
                       frag_num=max(1024, world), with_shadow=True)
    if world > 1 and opt == "adagrad":
        from swiftsnails_b200.parallel.cluster import attach_inbox
        attach_inbox(table, 2 * (2 * tps + ((tps + 127) // 128) * cfg.neg_pool), bf16_rows=True, barrier="flags")
    allk = torch.arange(rank, 2 * cfg.vocab, world, dtype=torch.int32, device=dev)
    for c in range(0, allk.numel(), 1 << 20):
        table.resolve(allk[c:c + (1 << 20)])
    torch.cuda.synchronize()
    grp.barrier()
    w = Word2VecWorker(cfg, table, dev)
    w.tma_mode = 33
    steps = 12
    toks = torch.from_numpy(synthetic_corpus(cfg.vocab, steps * cfg.tokens_per_step, seed=50 + rank).view(np.int32)).reshape(steps, -1).to(dev)
    s0 = w.sets[0]
    B = cfg.tokens_per_step
    names = ["keys+negsample", "dedup", "pull", "sgns_tc", "push"]
    acc = {n: 0.0 for n in names}
    push_parts = {"send": 0.0, "own": 0.0, "handoff": 0.0, "apply": 0.0}
    inbox = getattr(table, "inbox", None)
    if inbox is not None and os.environ.get("SSB200_SHIP_FIRST") == "0":
        inbox.ship_first = False
    if inbox is not None and os.environ.get("SSB200_OWN_STREAM") == "0":
        inbox.own_stream = None

    def phase(name, fn):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    from swiftsnails_b200.ops import sgns as sgns_ops
    for s in range(steps):
        t = toks[s]
        def keys():
            s0.keys_all[:B].copy_(t); torch.add(t, cfg.vocab, out=s0.keys_all[B:2 * B])
            w.sampler.sample(w.n_negs, out=s0.keys_all[2 * B:], key_offset=cfg.vocab)
        ts = [phase("k", keys),
              phase("d", lambda: s0.dedup(s0.keys_all, unique=s0.unique, inverse=s0.inverse, n_unique=s0.n_unique)),
              phase("p", lambda: table.pull(s0.unique, n_dev=s0.n_unique, rows_out=None, rows_bf16=s0.params_bf16, want_fp32=False,
                                            gslot_out=s0.gslot, grads_zero=s0.grads, n=w.n_keys)),
              phase("s", lambda: sgns_ops.sgns_block_tc(s0.params_bf16, s0.grads, s0.inverse[:B], s0.inverse[B:2 * B], cfg.sent_len,
                                                         cfg.window, cfg.seed, s * B, s0.inverse[2 * B:], 5.0, w.loss, mode=33))]
        if inbox is not None:
            inbox.marks = []
        ts.append(phase("u", lambda: w._push(s0)))
        grp.barrier()
        if s >= 2:
            for n, v in zip(names, ts):
                acc[n] += v
            if inbox is not None and inbox.marks:
                mk = inbox.marks
                for (n0, e0), (n1, e1) in zip(mk[:-1], mk[1:]):
                    push_parts[n1] += e0.elapsed_time(e1)
    out = [float(grp.all_reduce(acc[n] / (steps - 2), "max")) for n in names]
    parts = {k: round(float(grp.all_reduce(v / (steps - 2), "max")) * 1e3, 1) for k, v in push_parts.items()}
    if rank == 0:
        d = {n: round(float(v) * 1e3, 1) for n, v in zip(names, out)}
        d.update({"n_gpus": world, "optimizer": opt, "unit": "us per step (max over ranks)", "unique_keys": int(s0.n_unique.item()),
                  "sum_us": round(sum(out) * 1e3, 1), "push_parts_us": parts if inbox is not None else None})
        print(json.dumps(d))
    torch.cuda.synchronize()
    grp.barrier()
    grp.finish()
    grp.close()


if __name__ == "__main__":
    main()
