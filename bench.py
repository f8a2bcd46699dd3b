#!/usr/bin/env python
"""Headline benchmark: word2vec skip-gram negative-sampling training throughput (words/sec)
on the parameter server — BASELINE.json config 2 (synthetic Zipf corpus, vocab 1M, dim 128).

    python bench.py --gpus N --steps K --warmup W            # ours (N>1: launched by torchrun)
    python bench.py --impl reference --gpus N ...             # the reference's ZeroMQ CPU build

One rank per GPU hosts server shard g and worker g.  Weak scaling: every worker consumes
``--tokens-per-step`` tokens per step.  Timing: CUDA events around exactly K steps after W
warm-up steps, barrier + synchronize on both sides, max over ranks.  ``value`` is the
device-timed number with the token batches already resident on the device; ``e2e`` runs the
same K steps through the public ``Word2VecTrainer`` API including, every step, the H2D copy of
the step's token batch from pinned host memory and the D2H read of the step's loss.
The 1 GiB+ table is larger than L2 (126 MB), so no explicit L2 flush is used.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "word2vec_sgns_words_per_sec"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--tc-ctas", type=int, default=-1, help="persistent CTAs of the tensor-core kernel (-1: 148 on one GPU, "
                    "SSB200 default for several; 0: one per SM)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "nccl", "nccl_padded"],
                    help="nccl: all-to-all pull/push with exact splits (host sync per pull); nccl_padded: fixed buckets, no host sync, bf16 on the wire")
    ap.add_argument("--vocab", type=int, default=1_000_000)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--window", type=int, default=5)
    ap.add_argument("--negatives", type=int, default=5)
    ap.add_argument("--neg-pool", type=int, default=16)
    ap.add_argument("--tokens-per-step", type=int, default=524288,
                    help="minibatch per GPU; 524288 converges per token like 65536..262144 (benchmarks/batch_size_convergence.py)")
    ap.add_argument("--optimizer", default="adagrad", choices=["sgd", "adagrad"])
    ap.add_argument("--tile-impl", default="auto", choices=["auto", "cuda_core", "tcgen05", "tcgen05_direct"])
    ap.add_argument("--lr", type=float, default=0.05)
    ap.add_argument("--zipf-s", type=float, default=1.0)
    ap.add_argument("--shadow", type=int, default=1, help="bf16 mirror of the table rows, read by the pull (half the row bytes over NVLink); on at every GPU count so that the scaling curve is ONE configuration (it costs ~3%% on one GPU)")
    ap.add_argument("--inbox-bf16", type=int, default=1, help="ship inbox gradient rows as bf16")
    ap.add_argument("--inbox-barrier", default="flags", choices=["flags", "nccl", "chunks"], help="hand-off of the multi-GPU inbox push: flags / nccl = ship everything, then apply; chunks = 256-row chunks applied as they land (ship and apply overlap)")
    ap.add_argument("--inbox", type=int, default=1, help="multi-GPU AdaGrad push through owner inboxes (one-way) instead of remote atomics")
    ap.add_argument("--prepopulate", type=int, default=1, help="create all 2*vocab rows before timing (steady-state table)")
    ap.add_argument("--pipeline", type=int, default=1, help="overlap dedup+pull of step i+1 with SGNS+push of step i (2 streams)")
    ap.add_argument("--tc-mode", type=int, default=33, help="tcgen05 kernel mode bits (32 = cp.async gather warps (default), 1 = TMA gather4; 2 / 16 / 64 = perf experiments; 8 = round-1 serial kernel)")
    ap.add_argument("--slot-cache", type=int, default=1, help="worker-side key->slot cache (starts empty after the table is populated; fills from the warm-up steps on); on at every GPU count (same configuration for the whole scaling curve)")
    ap.add_argument("--windows", type=int, default=5, help="timed windows of exactly --steps steps each; the reported value is the median window")
    ap.add_argument("--sent-len", type=int, default=1024, help="tokens per sentence (pairs never cross sentences); divides tokens-per-step so the text arm reads whole minibatches")
    ap.add_argument("--text-arm", type=int, default=1, help="also time Word2VecTrainer.train_file on a text corpus file (the reference arm's input format)")
    ap.add_argument("--fp32-tile", type=int, default=1, help="also time a few steps with the fp32 CUDA-core tile kernel (config.fp32_tile_ms_per_step)")
    ap.add_argument("--ref-minibatch-lines", default="8,32", help="reference arm: minibatch sizes (sentences) to sweep; the fastest run that trains is reported")
    ap.add_argument("--ref-tokens", type=int, default=204_800, help="tokens per worker for the reference arm")
    return ap.parse_args()


DEFAULT_MULTI_GPU_TC_CTAS = 0      # persistent CTAs of the tensor-core kernel with several GPUs (0 = one per SM)


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def run_ours(args):
    import numpy as np
    import torch

    from swiftsnails_b200 import _lib
    from swiftsnails_b200.models.word2vec import (Word2VecConfig, Word2VecTrainer, Word2VecWorker,
                                                  synthetic_corpus)
    from swiftsnails_b200.parallel import group as ps_group
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.parallel.cluster import make_table
    from swiftsnails_b200.utils.clocks import ClockSampler

    rank, world, local = dist_env()
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun for --gpus > 1")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    # The job is run by the master, as in the reference: every rank registers with swift_master over ZeroMQ and gets
    # its rank, the route and the fragment map from it (parallel/group.py). No torch.distributed / NCCL on this
    # path; only the NCCL baseline arm (and the collective hand-off variant of the inbox) initialises it.
    nccl_arm = args.impl in ("nccl", "nccl_padded")
    grp = ps_group.init("torch" if (nccl_arm or args.inbox_barrier == "nccl") else "master", frag_num=max(1024, world))
    rank = grp.rank

    tile_impl = args.tile_impl
    if tile_impl == "auto":
        tile_impl = "tcgen05" if (_lib.has_symbol("ssb_sgns_block_tc") and args.dim == 128
                                  and os.environ.get("SSB200_NO_TC") != "1") else "cuda_core"
    cfg = Word2VecConfig(vocab=args.vocab, dim=args.dim, window=args.window, negatives=args.negatives,
                         neg_pool=args.neg_pool, tokens_per_step=args.tokens_per_step, lr=args.lr,
                         optimizer=args.optimizer, seed=1 + rank, tile_impl=tile_impl, zipf_s=args.zipf_s, sent_len=args.sent_len,
                         slot_cache=(args.impl == "ours" and bool(args.slot_cache)))
    # 2*vocab keys (input + output vectors) spread over `world` shards at load factor 0.5
    rows_per_shard = (2 * cfg.vocab + world - 1) // world
    use_inbox = args.impl == "ours" and args.inbox and world > 1 and cfg.optimizer == "adagrad"
    # bf16 mirror of the rows for the (bf16-only) pull; kept in step by owner-side writers, so it is
    # enabled when every push is applied by the owner (single GPU, or the inbox push)
    shadow_on = bool(args.shadow)
    with_shadow = (shadow_on or tile_impl == "tcgen05_direct") and args.impl == "ours" and tile_impl.startswith("tcgen05") and (world == 1 or use_inbox)
    table = make_table(rows_per_shard, cfg.dim, dev, with_state=(cfg.optimizer == "adagrad"),
                       init_scale=1.0 / cfg.dim, frag_num=max(1024, world), with_shadow=with_shadow)
    if use_inbox:
        from swiftsnails_b200.parallel.cluster import attach_inbox

        n_keys_max = 2 * cfg.tokens_per_step + ((cfg.tokens_per_step + 127) // 128) * cfg.neg_pool
        ib = attach_inbox(table, 2 * n_keys_max, bf16_rows=bool(args.inbox_bf16), barrier=args.inbox_barrier,
                          owner_pulls=(None if os.environ.get("SSB200_OWNER_PULLS") is None else os.environ["SSB200_OWNER_PULLS"] != "0"))
        if os.environ.get("SSB200_SHIP_FIRST") == "0" and hasattr(ib, "ship_first"):
            ib.ship_first = False
        if os.environ.get("SSB200_OWN_STREAM") == "0" and hasattr(ib, "own_stream"):
            ib.own_stream = None
    if nccl_arm:
        from swiftsnails_b200.parallel.nccl_baseline import NcclShardedTable

        table = NcclShardedTable(table, padded=(args.impl == "nccl_padded"), wire_bf16=(args.impl == "nccl_padded"))
    K, W, B = args.steps, args.warmup, cfg.tokens_per_step
    # the model is fully materialised before timing (every input and output vector exists, like a
    # dense model whose weights are allocated at init): rank r creates the keys == r (mod world)
    if args.prepopulate:
        all_keys = torch.arange(rank, 2 * cfg.vocab, world, dtype=torch.int32, device=dev)
        for c in range(0, all_keys.numel(), 1 << 20):
            table.pull(all_keys[c:c + (1 << 20)], want_fp32=False) if nccl_arm else table.resolve(all_keys[c:c + (1 << 20)])
        del all_keys
        torch.cuda.synchronize()
        grp.barrier()
    # two disjoint halves of the corpus: the device-timed arm and the end-to-end arm never see the same batch
    corpus = synthetic_corpus(cfg.vocab, 2 * (K + W) * B, s=cfg.zipf_s, seed=100 + rank)
    host_all = torch.from_numpy(corpus.view(np.int32)).reshape(2 * (K + W), B).pin_memory()
    host, host_e2e = host_all[:K + W], host_all[K + W:]

    def barrier():
        # this GPU idle, then every rank here (a host-side round through the master, ~0.1 ms of skew at most,
        # against windows of >= 10 ms; the reported time is the max over ranks)
        torch.cuda.synchronize()
        grp.barrier()

    def max_over_ranks(ms):
        return float(grp.all_reduce(float(ms), "max"))

    NW = max(1, args.windows)

    def median(xs):
        xs = sorted(xs)
        return xs[len(xs) // 2]

    # ---------------- device-timed arm: batches already on the device -------------------------
    # W warm-up steps, then NW windows of exactly K steps each, every window bracketed by a barrier + synchronize
    # and timed with CUDA events (max over ranks); the reported value is the median window. The K + W distinct
    # batches are cycled through (the table keeps training, so no window repeats another's state).
    worker = Word2VecWorker(cfg, table, dev, pipeline=bool(args.pipeline) and args.impl == "ours")
    worker.tma_mode = args.tc_mode
    tc_ctas = args.tc_ctas if args.tc_ctas >= 0 else (0 if world == 1 else DEFAULT_MULTI_GPU_TC_CTAS)
    if tile_impl.startswith("tcgen05"):
        _lib.call("ssb_sgns_tc_set_ctas", int(tc_ctas))
    dev_tokens = host.to(dev)
    nb = dev_tokens.shape[0]
    for s in range(W):
        worker.step(dev_tokens[s])
    worker.flush()
    barrier()
    sampler = ClockSampler(gpu_index=local).start()
    time.sleep(0.3)
    t0 = time.time()
    win_ms, launches, loss_dev, n_pairs = [], 0, None, 0.0
    cur = W
    for wdx in range(NW):
        worker.reset_loss()
        launches0 = _lib.launch_count
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for s in range(K):
            worker.step(dev_tokens[cur % nb])
            cur += 1
        worker.flush()
        e1.record()
        barrier()
        launches = _lib.launch_count - launches0
        win_ms.append(max_over_ranks(e0.elapsed_time(e1)))
        loss_dev = worker.mean_loss()
        n_pairs = float(worker.loss[1].item())
    t1 = time.time()
    ms = median(win_ms)
    n_unique_last = int(worker.n_unique.item())
    sampler.stop()
    clocks = sampler.summary(t0, t1)
    del dev_tokens

    # ---------------- end-to-end arm: public API, H2D tokens + D2H loss every step -------------
    trainer = Word2VecTrainer(cfg, table, dev, pipeline=bool(args.pipeline) and args.impl == "ours")
    trainer.worker.tma_mode = args.tc_mode
    trainer.worker.step_idx = worker.step_idx
    nbe = host_e2e.shape[0]
    for s in range(W):
        trainer.train_step(host_e2e[s])
    trainer.finish()
    e2e_ms, last_loss = [], None
    cur = W
    for wdx in range(NW):
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for s in range(K):
            trainer.train_step(host_e2e[cur % nbe])
            cur += 1
        last_loss = trainer.finish()
        trainer.worker.flush()
        f1.record()
        barrier()
        e2e_ms.append(max_over_ranks(f0.elapsed_time(f1)))
    ms_e2e = median(e2e_ms)

    # ---------------- text arm: the same public API fed from a corpus FILE (the reference arm's input) ----------
    # Word2VecTrainer.train_file: native parallel reader -> pinned batches -> H2D -> step -> D2H loss. The whole
    # call is timed (its first steps included), over K + W minibatches per rank.
    text = None
    if args.text_arm and args.impl == "ours" and B % args.sent_len == 0:
        import tempfile

        tdir = tempfile.mkdtemp(prefix="ssb200_text_")
        tpath = os.path.join(tdir, "corpus%d.txt" % rank)
        arr = host_e2e.numpy().view(np.uint32).reshape(-1, args.sent_len)
        with open(tpath, "w") as f:
            for row in arr:
                f.write(" ".join(map(str, row.tolist())))
                f.write("\n")
        barrier()
        g0, g1, gw = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        marks = {}

        def _mark(step):                      # after the W-th minibatch: start of the steady-state span
            if step == W:
                trainer.worker.flush()
                gw.record()
                marks["tokens_before"] = step * B
        trainer.on_step = _mark
        tw0 = time.time()
        g0.record()
        # negatives stay on the synthetic corpus' noise distribution, as in the other arms
        n_tok_text, text_loss = trainer.train_file(tpath, corpus_noise=False)
        trainer.worker.flush()
        g1.record()
        barrier()
        trainer.on_step = None
        ms_text = max_over_ranks(g0.elapsed_time(g1))
        ms_steady = max_over_ranks(gw.elapsed_time(g1)) if marks else ms_text
        n_steady = trainer.steps_done - (W if marks else 0)
        text = {"value": float(n_steady) * B * world / (ms_steady / 1e3), "unit": "words/s", "steps": n_steady,
                "ms_per_step": ms_steady / max(n_steady, 1),
                "whole_call": {"steps": trainer.steps_done, "ms": ms_text, "wall_ms": (time.time() - tw0) * 1e3,
                               "note": "includes opening, mapping and indexing the file and the first (warm-up) steps"},
                "file_bytes_per_rank": os.path.getsize(tpath), "loss_last": text_loss, "host_cpus": os.cpu_count(),
                "includes": "text parse by the native loader threads + H2D + step + D2H loss, every step"}
        os.unlink(tpath)
        os.rmdir(tdir)

    # ---------------- fp32 tile kernel (CUDA cores, fp32 operands): the step at the reference's operand precision ----
    fp32_ms = None
    if args.fp32_tile and args.impl == "ours" and tile_impl.startswith("tcgen05"):
        import dataclasses

        cfg32 = dataclasses.replace(cfg, tile_impl="cuda_core")
        w32 = Word2VecWorker(cfg32, table, dev, pipeline=bool(args.pipeline))
        w32.step_idx = trainer.worker.step_idx
        n32 = min(8, host.shape[0])
        d32 = host[:n32].to(dev)
        for s in range(3):
            w32.step(d32[s % n32])
        w32.flush()
        barrier()
        h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h0.record()
        for s in range(3, 8):
            w32.step(d32[s % n32])
        w32.flush()
        h1.record()
        barrier()
        fp32_ms = max_over_ranks(h0.elapsed_time(h1)) / 5.0
        del w32, d32

    # ---------------- lost updates: table full / inbox overflow must both be zero -----------------------------
    health = {"table_full": int(table.stats[1].item()) if hasattr(table, "stats") else 0, "inbox_overflow": 0}
    if getattr(table, "inbox", None) is not None:
        health["inbox_overflow"] = int(table.inbox.overflow.item())
        if hasattr(table.inbox, "err"):
            health["inbox_overflow"] += int(table.inbox.err.item())      # an apply wait ran out: rows were skipped
    if world > 1:
        health = {k: int(grp.all_reduce(v, "sum")) for k, v in health.items()}

    words = float(K) * B * world
    value = words / (ms / 1e3)
    e2e_value = words / (ms_e2e / 1e3)
    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "words/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if tile_impl.startswith("tcgen05") else "fp32",
            "data": "synthetic Zipf(s=1) corpus, random-init table", "impl": args.impl,
            "config": {"model": "word2vec skip-gram negative sampling", "vocab": cfg.vocab, "dim": cfg.dim,
                       "window": cfg.window, "negatives": cfg.negatives,
                       "neg_pool_per_128_pair_tile": cfg.neg_pool, "global_batch": B * world,
                       "tokens_per_step_per_gpu": B, "seq_len": cfg.sent_len, "optimizer": cfg.optimizer,
                       "parallelism": "ps%d (1 server shard + 1 worker per GPU, %s pull/push)" % (world, "NCCL all-to-all" if nccl_arm else "P2P"),
                       "bf16_row_mirror": with_shadow, "worker_slot_cache": bool(cfg.slot_cache),
                       "inbox_push": bool(use_inbox), "inbox_rows_bf16": bool(use_inbox and args.inbox_bf16),
                       "inbox_handoff": args.inbox_barrier if use_inbox else None,
                       "inbox_rows_cross_nvlink_as": ("loads by the owner's apply kernel (senders fill local outboxes)" if getattr(getattr(table, "inbox", None), "owner_pulls", False) else "stores by the sender") if use_inbox else None, "tc_kernel_ctas": tc_ctas or 148,
                       "control_plane": {"single": "one process", "master": "swift_master over ZeroMQ (registration, route, fragment map, barriers); no torch.distributed / NCCL",
                                         "torch": "torch.distributed (NCCL)"}[grp.kind],
                       "same_switches_at_every_gpu_count": "slot cache, bf16 row mirror and bf16 inbox rows are on for every N; the inbox itself only exists with peers (N > 1)",
                       "timed_windows": NW, "window_ms_per_step": [round(x / K, 5) for x in win_ms],
                       "e2e_window_ms_per_step": [round(x / K, 5) for x in e2e_ms],
                       "fp32_tile_ms_per_step": fp32_ms, "e2e_text": text,
                       "table": "all 2*vocab rows created before timing" if args.prepopulate else "empty at start, rows created lazily inside the timed region",
                       "pipeline": "3 streams: dedup+pull of step i+1 | SGNS of step i | push of step i-1 (staleness <= 1 step)" if worker.pipeline else "off",
                       "tile_impl": tile_impl, "pairs_per_step_per_gpu": n_pairs / K, "unique_keys_last_step": n_unique_last,
                       "precision": ("bf16 tensor-core operands (tcgen05), fp32 accumulate, fp32 master table / AdaGrad state, fp32 gradient accumulation on the worker"
                                     + ("; gradient rows cross NVLink as bf16 and are applied in fp32 by the owner" if (use_inbox and args.inbox_bf16) else "")),
                       "l2": "no flush: table (%.2f GiB/shard) and caches exceed the 126 MB L2"
                             % (table.owned[0].nbytes() / 2 ** 30)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "words/s", "ms_per_step": ms_e2e / K,
                    "h2d_bytes_per_step": trainer.h2d_bytes, "d2h_bytes_per_step": trainer.d2h_bytes},
            "gpu_launches": launches,
            "loss": {"device_arm_mean": loss_dev, "e2e_last": last_loss},
            "health": health,
        }
        print(json.dumps(out))
    if world > 1:
        torch.cuda.synchronize()
        grp.barrier()
        if grp.kind == "torch":
            import torch.distributed as dist

            dist.destroy_process_group()
        else:
            grp.finish()          # WORKER_FINISH_WORK -> SERVER_TOLD_TO_TERMINATE -> ack (no dump in the benchmark)
            grp.close()
    if health["table_full"] or health["inbox_overflow"]:
        raise SystemExit("bench.py: updates were lost during the run: %r" % (health,))


def run_reference(args):
    try:
        from baseline.run_reference import run as ref_run
    except Exception as e:  # harness missing
        print(json.dumps({"impl": "reference", "unavailable": "reference harness not importable: %r" % (e,)}))
        return
    rank, world, _ = dist_env()
    if rank != 0:
        return  # the reference arm is CPU-only: rank 0 launches all master/server/worker processes
    print(json.dumps(ref_run(args)))


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
