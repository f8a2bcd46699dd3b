"""Train word2vec on the GPU parameter server from the command line — the GPU counterpart of
``swift_worker -config worker.conf -data data.txt`` (/root/reference/src/tools/run_worker.sh:1-5):

    python -m swiftsnails_b200.train -config worker.conf -data corpus.txt -out params.txt
    tools/launch_gpu_job.sh 8 -m swiftsnails_b200.train -config worker.conf -data corpus.txt -out params.txt

Same config grammar and key names as the reference's apps read (``dim``, ``vocab``, ``window``,
``negative``, ``learning_rate``, ``num_iters``) plus the optional keys of this
framework (``update: sgd|adagrad``, ``table_seed``, ``init_scale``, ``tokens_per_step``, ``sent_len``,
``param_dump_path``), and the server's ``param_backup_period`` / ``param_backup_root``: every that many
minibatches each rank writes ``<root>/param-<rank>-<n>.txt`` (system/server/init.h:128-149).  Same data format (one sentence of integer word ids per line) and the same
``key \\t value`` text dump at the end — one file per rank (``<out>.<rank>`` when there are several),
which is what the reference's servers produce (one dump per server, server/terminate.h:36).

One process per GPU hosts server shard g and worker g; worker g trains on the lines whose index is
g modulo the number of GPUs (the reference gets that split from Hadoop streaming), and every worker
runs the same number of minibatches.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import time


def parse_args(argv=None):
    ap = argparse.ArgumentParser(prog="swiftsnails_b200.train", description=__doc__.split("\n\n")[0])
    ap.add_argument("-config", "--config", required=True, help="config file (key: value, `import`, # comments)")
    ap.add_argument("-data", "--data", required=True, help="corpus: one sentence of integer word ids per line")
    ap.add_argument("-out", "--out", default=None, help="parameter dump (text, reference format); default: config key param_dump_path")
    ap.add_argument("-app", "--app", default="word2vec", choices=["word2vec"])
    ap.add_argument("--epochs", type=int, default=None, help="passes over the data (default: config key num_iters)")
    ap.add_argument("--log-every", type=int, default=0)
    ap.add_argument("--binary", action="store_true", help="write the raw binary shard dump instead of text")
    ap.add_argument("--dump-state", action="store_true",
                    help="text dump carries the optimizer state too (Param-style lines, 9 digits): exact resume")
    ap.add_argument("--resume", default=None, help="load this dump (text or binary; <path>.<rank> with several GPUs) before training")
    return ap.parse_args(argv)


def word2vec_config(conf, world=1):
    """Config file -> Word2VecConfig (pure host code: testable without a GPU)."""
    from .models.word2vec import Word2VecConfig

    dim = conf.to_int32("dim", 128)
    window = conf.to_int32("window", 5)
    sent_len = conf.to_int32("sent_len", 1000)
    tps = conf.to_int32("tokens_per_step", 524288)
    tps = max(sent_len, tps // sent_len * sent_len)          # whole sentences per minibatch
    update = conf.get("update", "adagrad").strip().lower()
    if update not in ("sgd", "adagrad"):
        raise ValueError("update must be sgd or adagrad, not %r" % update)
    return Word2VecConfig(
        vocab=int(conf.get("vocab", "1000000")), dim=dim, window=window, negatives=conf.to_int32("negative", 5),
        sent_len=sent_len, tokens_per_step=tps, lr=conf.to_float("learning_rate", 0.05), optimizer=update,
        seed=conf.to_int32("table_seed", 1),
        # the tensor-core kernel needs 128-wide rows and a window of at most 8; anything else runs on the CUDA cores
        tile_impl="tcgen05" if (dim == 128 and window <= 8) else "cuda_core")


def count_sentences(path):
    """Lines the loader will turn into a sentence: the first non-blank character is a digit (the same rule as
    the native reader, csrc/core/corpus_loader.h: corpus_is_sentence — it skips, and does not number, anything
    else, so every rank derives the same number of minibatches from this count)."""
    n = 0
    with open(path, "rb") as f:
        for line in f:
            t = line.lstrip(b" \t\r")
            if t[:1].isdigit():
                n += 1
    return n


def total_steps(n_sentences, world, cfg, epochs=1):
    """Minibatches EVERY rank can run: worker i reads the lines with index % world == i, so the smallest
    share is floor(n / world) sentences per pass; passes are read back to back (a minibatch may span
    the end of one pass and the start of the next), the last minibatch may be partly padding."""
    per_rank = n_sentences // world
    per_batch = cfg.tokens_per_step // cfg.sent_len
    return math.ceil(epochs * per_rank / per_batch) if per_rank > 0 else 0


def main(argv=None):
    args = parse_args(argv)
    from .utils.config import ConfigParser

    conf = ConfigParser(args.config)
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cfg = word2vec_config(conf, world)
    epochs = args.epochs if args.epochs is not None else conf.to_int32("num_iters", 1)
    out = args.out or conf.get("param_dump_path", "")

    import torch
    import torch.distributed as dist

    from .models.word2vec import Word2VecTrainer
    from .parallel.cluster import attach_inbox, make_table
    from .utils import dump

    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    adagrad = cfg.optimizer == "adagrad"
    init_scale = conf.to_float("init_scale", 1.0 / cfg.dim)
    table = make_table((2 * cfg.vocab + world - 1) // world, cfg.dim, dev, with_state=adagrad, init_scale=init_scale,
                       frag_num=max(conf.to_int32("frag_num", 1024), world), seed=cfg.seed,
                       with_shadow=(world >= 4 and cfg.tile_impl == "tcgen05" and adagrad))
    if world > 1 and adagrad:
        n_keys_max = 2 * cfg.tokens_per_step + ((cfg.tokens_per_step + 127) // 128) * cfg.neg_pool
        attach_inbox(table, 2 * n_keys_max, bf16_rows=True, barrier="flags")
    n_sent = count_sentences(args.data)
    steps = total_steps(n_sent, world, cfg, epochs)
    if steps == 0:
        raise SystemExit("corpus %s has fewer sentences (%d) than workers (%d)" % (args.data, n_sent, world))
    trainer = Word2VecTrainer(cfg, table, dev, pipeline=(cfg.tile_impl == "tcgen05"))
    if args.resume:
        rpath = args.resume if world == 1 else "%s.%d" % (args.resume, rank)
        with open(rpath, "rb") as f:
            is_binary = f.read(4) == b"SSB2"
        loaded = (dump.load_table_binary if is_binary else dump.load_table)(table, rpath)
        if rank == 0:
            print("resumed %d rows from %s" % (loaded, rpath))
    # periodic parameter backup, as the reference's server does every param_backup_period push requests
    # (system/server/init.h:128-149); here a push request is a minibatch of this rank's worker
    from .framework import SwiftServer

    server = SwiftServer(table, conf)
    if server.backup_period > 0:
        os.makedirs(server.backup_root, exist_ok=True)

        def _backup(step):
            trainer.worker.flush()
            server.on_push()
        trainer.on_step = _backup
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.time()
    tokens, loss = trainer.train_file(args.data, epochs=epochs, log_every=args.log_every if rank == 0 else 0,
                                      split=(world, rank) if world > 1 else None, max_steps=steps)
    trainer.worker.flush()
    torch.cuda.synchronize()
    if world > 1:
        # inbox pushes hand off in lock step: a rank whose share ran out early (fewer sentences than the
        # others) would leave them waiting, which is why everyone runs exactly `steps` minibatches
        assert trainer.steps_done == steps, (trainer.steps_done, steps)
        t = torch.tensor([float(tokens)], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        tokens_all = float(t.item())
        dist.barrier()
    else:
        tokens_all = float(tokens)
    dt = time.time() - t0
    rows = 0
    if out:
        path = out if world == 1 else "%s.%d" % (out, rank)
        if args.binary:
            rows = dump.save_table_binary(table, path)
        elif args.dump_state:
            rows = dump.save_table(table, path, with_state=True, digits=9)
        else:
            rows = dump.save_table(table, path)
    if world > 1:
        dist.barrier()
    if rank == 0:
        print("RESULT " + json.dumps({"app": "word2vec", "gpus": world, "tokens": tokens_all, "seconds": round(dt, 3),
                                      "words_per_s": round(tokens_all / max(dt, 1e-9), 1), "loss_per_pair": loss,
                                      "steps_per_rank": steps, "dump": out or None, "rows_dumped_rank0": rows,
                                      "ids_out_of_vocab_rank0": trainer.bad_tokens, "health_rank0": trainer.health(),
                                      "backups_rank0": server.push_counter // server.backup_period if server.backup_period > 0 else 0}))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
