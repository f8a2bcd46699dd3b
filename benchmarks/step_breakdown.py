#!/usr/bin/env python
"""Per-phase device time of the word2vec step (CUDA events around each op, non-pipelined), at the
launched world size: key build + negative sampling, dedup, pull, SGNS tiles, push.  Max over ranks."""
import json, os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker, synthetic_corpus
    from swiftsnails_b200.parallel.cluster import make_table

    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local); torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    opt = sys.argv[1] if len(sys.argv) > 1 else "adagrad"
    tps = int(sys.argv[2]) if len(sys.argv) > 2 else 524288
    cfg = Word2VecConfig(tile_impl="tcgen05", optimizer=opt, lr=0.05, tokens_per_step=tps, seed=1 + rank)
    # same table options as bench.py: bf16 row mirror from 4 GPUs, inbox push (bf16 rows, flag hand-off) from 2
    table = make_table((2 * cfg.vocab + world - 1) // world, cfg.dim, dev, with_state=(opt == "adagrad"), init_scale=1.0 / cfg.dim,
                       frag_num=max(1024, world), with_shadow=(world >= 4))
    if world > 1 and opt == "adagrad":
        from swiftsnails_b200.parallel.cluster import attach_inbox
        attach_inbox(table, 2 * (2 * tps + ((tps + 127) // 128) * cfg.neg_pool), bf16_rows=True, barrier="flags")
    allk = torch.arange(rank, 2 * cfg.vocab, world, dtype=torch.int32, device=dev)
    for c in range(0, allk.numel(), 1 << 20):
        table.resolve(allk[c:c + (1 << 20)])
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    w = Word2VecWorker(cfg, table, dev)
    steps = 12
    toks = torch.from_numpy(synthetic_corpus(cfg.vocab, steps * cfg.tokens_per_step, seed=50 + rank).view(np.int32)).reshape(steps, -1).to(dev)
    s0 = w.sets[0]
    B = cfg.tokens_per_step
    names = ["keys+negsample", "dedup", "pull", "sgns_tc", "push"]
    acc = {n: 0.0 for n in names}

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
                                                         cfg.window, cfg.seed, s * B, s0.inverse[2 * B:], 5.0, w.loss, mode=1)),
              phase("u", lambda: w._push(s0))]
        if world > 1: dist.barrier()
        if s >= 2:
            for n, v in zip(names, ts):
                acc[n] += v
    out = torch.tensor([acc[n] / (steps - 2) for n in names], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(out, op=dist.ReduceOp.MAX)
    if rank == 0:
        d = {n: round(float(v) * 1e3, 1) for n, v in zip(names, out)}
        d.update({"n_gpus": world, "optimizer": opt, "unit": "us per step (max over ranks)", "unique_keys": int(s0.n_unique.item()),
                  "sum_us": round(float(out.sum()) * 1e3, 1)})
        print(json.dumps(d))
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
