"""Train word2vec or sparse logistic regression on the GPU parameter server from the command line — the GPU
counterpart of ``swift_worker -config worker.conf -data data.txt`` (/root/reference/src/tools/run_worker.sh:1-5;
the apps the reference ships are named at src/tools/copy_exec.sh:4-9):

    python -m swiftsnails_b200.train -config worker.conf -data corpus.txt -out params.txt
    python -m swiftsnails_b200.train -app lr -config lr.conf -data examples.txt -out weights.txt
    tools/launch_gpu_job.sh 8 -m swiftsnails_b200.train -config worker.conf -data corpus.txt -out params.txt

With several GPUs the job is run by the MASTER, as in the reference: every rank registers with ``swift_master`` over
ZeroMQ (rank 0 starts one unless ``master_addr`` in the config names a running one), gets its rank, the route and the
fragment map from it, reports WORKER_FINISH_WORK at the end and dumps its shard when told SERVER_TOLD_TO_TERMINATE
(parallel/group.py). torch.distributed / NCCL is not initialised.

``-app lr`` reads ``label key:value key:value …`` lines (value defaults to 1; the CPU worker's format,
csrc/apps/algorithms.h) and the keys ``feature_bits`` (hashed feature space 2^bits; keys are taken modulo it),
``max_features`` per example, ``minibatch`` (examples per step), ``update: sgd|adagrad|ftrl``, ``learning_rate``,
``l1`` / ``l2`` (FTRL), ``num_iters``.

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
    ap.add_argument("-app", "--app", default="word2vec", choices=["word2vec", "lr"])
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


def _start(args):
    """Config, group, device — shared by the apps."""
    from .parallel import group as ps_group
    from .utils.config import ConfigParser

    conf = ConfigParser(args.config)
    local = int(os.environ.get("LOCAL_RANK", "0"))
    import torch

    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    master = conf.get("master_addr", "").strip() or None
    grp = ps_group.init("master", master_addr=master, frag_num=conf.to_int32("frag_num", 1024))
    return conf, grp, dev


def main_lr(args):
    """Sparse logistic regression: fused GPU step (csrc/kernels/lr.cu) over a `label key:value …` file."""
    import ctypes as C

    import numpy as np
    import torch

    from . import _lib
    from .framework import PushMethod, SwiftServer
    from .models.logistic_regression import LRConfig, LRWorker
    from .ops._util import u32_tensor
    from .parallel.cluster import make_table
    from .utils import dump

    conf, grp, dev = _start(args)
    rank, world = grp.rank, grp.world
    update = conf.get("update", "adagrad").strip().lower()
    bits = conf.to_int32("feature_bits", 24)
    max_feat = conf.to_int32("max_features", 64)
    epochs = args.epochs if args.epochs is not None else conf.to_int32("num_iters", 1)
    out = args.out or conf.get("param_dump_path", "")
    method = PushMethod(update, lr=conf.to_float("learning_rate", 0.05))          # resolves the rule: state words, init kind
    lib = _lib.core()
    lib.ssb_lr_parse_file.restype = C.c_longlong
    lib.ssb_lr_parse_file.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_ulonglong,
                                      C.c_int, C.c_int, C.c_void_p]
    stats = (C.c_longlong * 2)()
    n_all = lib.ssb_lr_parse_file(args.data.encode(), max_feat, None, None, None, 0, 1 << bits, 1, 0, stats)
    if n_all < 0:
        raise SystemExit("cannot read %s" % args.data)
    n_mine = n_all // world                       # every rank runs the same number of steps
    if n_mine == 0:
        raise SystemExit("%s has fewer examples (%d) than workers (%d)" % (args.data, n_all, world))
    keys = np.empty((n_mine, max_feat), dtype=np.uint32)
    vals = np.empty((n_mine, max_feat), dtype=np.float32)
    labels = np.empty(n_mine, dtype=np.float32)
    got = lib.ssb_lr_parse_file(args.data.encode(), max_feat, keys.ctypes.data, vals.ctypes.data, labels.ctypes.data, n_mine,
                                1 << bits, world, rank, stats)
    assert got == n_mine, (got, n_mine)
    batch = min(conf.to_int32("minibatch", 65536), n_mine)
    cfg = LRConfig(feature_bits=bits, n_numeric=0, n_categorical=max_feat, batch_size=batch, lr=method.hyper["lr"],
                   optimizer=update, l1=conf.to_float("l1", 0.0), l2=conf.to_float("l2", 0.0), fused=True)
    rows = conf.to_int32("table_rows", min(1 << bits, 1 << 26))
    table = make_table((rows + world - 1) // world, 1, dev, with_state=(method.state_words or False),
                       init_scale=conf.to_float("init_scale", 0.0), init_kind=method.init_kind,
                       frag_num=max(conf.to_int32("frag_num", 1024), world), seed=conf.to_int32("table_seed", 1))
    if args.resume:
        rpath = args.resume if world == 1 else "%s.%d" % (args.resume, rank)
        with open(rpath, "rb") as f:
            is_binary = f.read(4) == b"SSB2"
        (dump.load_table_binary if is_binary else dump.load_table)(table, rpath)
    wk = LRWorker(cfg, table, dev)
    server = SwiftServer(table, conf)
    if server.backup_period > 0:
        os.makedirs(server.backup_root, exist_ok=True)
    kd, vd, yd = u32_tensor(keys, dev), torch.from_numpy(vals).to(dev), torch.from_numpy(labels).to(dev)
    steps_per_epoch = n_mine // batch
    torch.cuda.synchronize()
    grp.barrier()
    t0 = time.time()
    first = last = None
    for ep in range(epochs):
        for s in range(steps_per_epoch):
            lo = s * batch
            wk.reset_loss()
            wk.step(kd[lo:lo + batch], vd[lo:lo + batch], yd[lo:lo + batch])
            if server.backup_period > 0:
                server.on_push()
            if (args.log_every and rank == 0 and s % args.log_every == 0) or (s == steps_per_epoch - 1) or (ep == 0 and s == 0):
                last = wk.mean_loss()
                first = last if first is None else first
                if args.log_every and rank == 0:
                    print("epoch %d step %d loss %.5f" % (ep, s, last))
    torch.cuda.synchronize()
    dt = time.time() - t0
    examples = float(grp.all_reduce(epochs * steps_per_epoch * batch, "sum"))
    health = wk.health()
    rows_out = [0]

    def _dump():
        if out:
            path = out if world == 1 else "%s.%d" % (out, rank)
            if args.binary:
                rows_out[0] = dump.save_table_binary(table, path)
            else:
                rows_out[0] = dump.save_table(table, path, with_state=args.dump_state, digits=9 if args.dump_state else 6)

    grp.finish(dump=_dump)        # WORKER_FINISH_WORK -> SERVER_TOLD_TO_TERMINATE -> dump -> ack (one process: just the dump)
    if rank == 0:
        print("RESULT " + json.dumps({"app": "lr", "gpus": world, "control_plane": grp.kind, "examples": examples,
                                      "seconds": round(dt, 3), "examples_per_s": round(examples / max(dt, 1e-9), 1),
                                      "loss_first": first, "loss_last": last, "update": update,
                                      "features_dropped_rank0": int(stats[1]), "health_rank0": health, "dump": out or None,
                                      "rows_dumped_rank0": rows_out[0]}))
    grp.close()
    if health.get("table_full") or health.get("touched_overflow"):
        raise SystemExit("updates were lost: %r" % (health,))
    return 0


def main(argv=None):
    args = parse_args(argv)
    if args.app == "lr":
        return main_lr(args)
    conf, grp, dev = _start(args)
    rank, world = grp.rank, grp.world
    cfg = word2vec_config(conf, world)
    epochs = args.epochs if args.epochs is not None else conf.to_int32("num_iters", 1)
    out = args.out or conf.get("param_dump_path", "")

    import torch

    from .models.word2vec import Word2VecTrainer
    from .parallel.cluster import attach_inbox, make_table
    from .utils import dump

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
    grp.barrier()
    t0 = time.time()
    tokens, loss = trainer.train_file(args.data, epochs=epochs, log_every=args.log_every if rank == 0 else 0,
                                      split=(world, rank) if world > 1 else None, max_steps=steps)
    trainer.worker.flush()
    torch.cuda.synchronize()
    if world > 1:
        # inbox pushes hand off in lock step: a rank whose share ran out early (fewer sentences than the
        # others) would leave them waiting, which is why everyone runs exactly `steps` minibatches
        assert trainer.steps_done == steps, (trainer.steps_done, steps)
    tokens_all = float(grp.all_reduce(float(tokens), "sum"))
    dt = time.time() - t0
    rows = [0]

    def _dump():
        # runs when the master says SERVER_TOLD_TO_TERMINATE, i.e. after EVERY worker has reported
        # WORKER_FINISH_WORK: no rank is still pushing into this shard (server/terminate.h:27-45)
        if out:
            path = out if world == 1 else "%s.%d" % (out, rank)
            if args.binary:
                rows[0] = dump.save_table_binary(table, path)
            elif args.dump_state:
                rows[0] = dump.save_table(table, path, with_state=True, digits=9)
            else:
                rows[0] = dump.save_table(table, path)

    health = trainer.health()
    grp.finish(dump=_dump)        # WORKER_FINISH_WORK -> SERVER_TOLD_TO_TERMINATE -> dump -> ack (one process: just the dump)
    if rank == 0:
        print("RESULT " + json.dumps({"app": "word2vec", "gpus": world, "control_plane": grp.kind, "tokens": tokens_all,
                                      "seconds": round(dt, 3),
                                      "words_per_s": round(tokens_all / max(dt, 1e-9), 1), "loss_per_pair": loss,
                                      "steps_per_rank": steps, "dump": out or None, "rows_dumped_rank0": rows[0],
                                      "ids_out_of_vocab_rank0": trainer.bad_tokens, "health_rank0": health,
                                      "backups_rank0": server.push_counter // server.backup_period if server.backup_period > 0 else 0}))
    grp.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
