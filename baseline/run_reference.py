"""Reference arm of bench.py: the UNMODIFIED FrankSzn/SwiftSnails headers (baseline/_ref/src, a
byte-identical copy of /root/reference/src) built offline with the shim headers in
baseline/shim and run through the reference's own public API and stock ZeroMQ code path:
1 master + N servers + N workers on localhost CPUs (N = --gpus, so the topology matches ours:
one server shard + one worker per "slot").  Same metric and config as the main arm: word2vec
skip-gram negative sampling, synthetic Zipf(s=1) corpus, vocab 1M, dim 128, window 5, 5 negatives.

The reference's own apps are absent from its snapshot; baseline/ref_apps/ref_ps.cpp supplies the
application layer against the reference API.  The fixed sleeps in the reference (3 s before
training, SwiftWorker.h:105; 5 s before termination, worker/terminate.h:39) are outside the
timed region (words / time inside train()).
"""
from __future__ import annotations

import glob
import json
import os
import shutil
import socket
import subprocess
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
BIN = os.path.join(REF, "bin", "ref_ps")


def _libzmq():
    for sp in sys.path:
        hits = glob.glob(os.path.join(sp, "pyzmq.libs", "libzmq*.so*"))
        if hits:
            return hits[0]
    return None


def ensure_built():
    """Copy the reference (if the copy is missing) and compile ref_ps. Returns None or an error."""
    src = os.path.join(REF, "src")
    if not os.path.isdir(src):
        if not os.path.isdir("/root/reference/src"):
            return "baseline/_ref/src missing and /root/reference not available"
        os.makedirs(REF, exist_ok=True)
        shutil.copytree("/root/reference/src", src)
    app = os.path.join(HERE, "ref_apps", "ref_ps.cpp")
    if os.path.exists(BIN) and os.path.getmtime(BIN) >= os.path.getmtime(app):
        return None
    lz = _libzmq()
    if lz is None:
        return "libzmq (pyzmq.libs) not found"
    os.makedirs(os.path.dirname(BIN), exist_ok=True)
    cmd = ["g++", "-std=c++11", "-O2", "-pthread", "-w", "-I", os.path.join(HERE, "shim"), "-I", src, app, "-o", BIN,
           "-L" + os.path.dirname(lz), "-l:" + os.path.basename(lz), "-Wl,-rpath," + os.path.dirname(lz)]
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        return "reference build failed: " + p.stderr.strip().splitlines()[-1][:200]
    return None


ORACLE = os.path.join(REF, "bin", "ref_oracle")


def ensure_oracle_built():
    """Compile baseline/ref_apps/ref_oracle.cpp (the reference's own hashfrag / sparse table as a test oracle)
    against the unmodified reference copy. Returns None or an error string."""
    err = ensure_built()
    if err:
        return err
    app = os.path.join(HERE, "ref_apps", "ref_oracle.cpp")
    if os.path.exists(ORACLE) and os.path.getmtime(ORACLE) >= os.path.getmtime(app):
        return None
    lz = _libzmq()
    cmd = ["g++", "-std=c++11", "-O1", "-pthread", "-w", "-I", os.path.join(HERE, "shim"), "-I", os.path.join(REF, "src"), app, "-o", ORACLE,
           "-L" + os.path.dirname(lz), "-l:" + os.path.basename(lz), "-Wl,-rpath," + os.path.dirname(lz)]
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        return "oracle build failed: " + p.stderr.strip()[-400:]
    return None


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _write_conf(wd, n_nodes, extra):
    port = _free_port()
    common = {
        "async_exec_num": 4, "listen_thread_num": 1, "init_timeout": 120, "master_time_out": 120,
        "master_longest_alive_duration": 3600, "frag_num": 1024, "shard_num": 8, "param_backup_period": 0,
        "param_backup_root": wd, "num_iters": 1, "learning_rate": 0.05, "async_channel_thread_num": 2, "optimizer": "adagrad", "neg_pool": 16, "prepopulate": 0,
        "local_train": 0, "master_addr": "tcp://127.0.0.1:%d" % port,
    }
    common.update(extra)
    with open(os.path.join(wd, "common.conf"), "w") as f:
        for k, v in common.items():
            f.write("%s: %s\n" % (k, v))
    with open(os.path.join(wd, "master.conf"), "w") as f:
        f.write("import common.conf\nlisten_addr: tcp://127.0.0.1:%d\nexpected_node_num: %d\n" % (port, n_nodes))
    with open(os.path.join(wd, "node.conf"), "w") as f:
        f.write("import common.conf\nlisten_addr:\n")


def launch(n, app, conf_extra, data_files=None, timeout=1800):
    wd = tempfile.mkdtemp(prefix="ssb200_ref_")
    _write_conf(wd, 2 * n, conf_extra)
    env = dict(os.environ)

    def spawn(role, tag, extra=()):
        so = open(os.path.join(wd, tag + ".out"), "w")
        se = open(os.path.join(wd, tag + ".err"), "w")
        conf = "master.conf" if role == "master" else "node.conf"
        p = subprocess.Popen([BIN, "-role", role, "-config", conf, "-app", app] + list(extra), cwd=wd, stdout=so, stderr=se, env=env)
        p._files = (so, se)
        return p

    procs = [spawn("master", "master")]
    time.sleep(0.3)
    servers = [spawn("server", "server%d" % i) for i in range(n)]
    workers = [spawn("worker", "worker%d" % i, ["-data", data_files[i]] if data_files else []) for i in range(n)]
    deadline = time.time() + timeout
    outs = []
    failed = None
    for p in workers:
        try:
            p.wait(timeout=max(1.0, deadline - time.time()))
        except subprocess.TimeoutExpired:
            failed = "reference worker timed out"
            break
    for p in workers + servers + procs:
        if p.poll() is None:
            try:
                p.wait(timeout=30 if failed is None else 1)
            except subprocess.TimeoutExpired:
                p.kill()
        for f in p._files:
            f.close()
    for i in range(n):
        with open(os.path.join(wd, "worker%d.out" % i)) as f:
            outs.append(f.read())
    errs = ""
    if failed or any(w.returncode != 0 for w in workers):
        with open(os.path.join(wd, "worker0.err")) as f:
            errs = f.read()[-400:]
        failed = failed or "reference worker exited %s: %s" % ([w.returncode for w in workers], errs.replace("\n", " | "))
    return outs, failed, wd


def _parse(stdout, tag):
    for line in stdout.splitlines():
        if line.startswith(tag + " "):
            t = line.split()[1:]
            return {t[i]: float(t[i + 1]) for i in range(0, len(t) - 1, 2)}
    return None


def _threads_per_worker(n):
    """Worker compute threads (the reference's async_channel_thread_num): what the host can give each of the
    n workers next to n servers (4 handler threads each), between 1 and 8."""
    cpus = os.cpu_count() or 8
    return max(1, min(8, (cpus - 1) // (2 * n)))


def _one_run(n, conf, data_files):
    outs, failed, _ = launch(n, "word2vec", conf, data_files)
    if failed:
        return None, failed
    res = [_parse(o, "RESULT") for o in outs]
    if any(r is None for r in res):
        return None, "reference worker produced no RESULT line"
    return res, None


def run(args):
    """Returns the JSON dict for `bench.py --impl reference`.

    Same workload as the main arm: SGNS over tiles of 128 consecutive centre tokens that share `neg_pool`
    negatives (weight m_i*K/neg_pool), window 5 with shrink, AdaGrad (lr and eps of the main arm) applied by the
    server, gradients summed per minibatch and pushed once per key, rows created lazily by the first pull.
    The minibatch size is swept over `--ref-minibatch-lines` and the fastest run is reported; the run is
    rejected unless its loss is finite, not above the untrained value ln2*(1+K) (2 % slack) and lower at the end than at the start."""
    import math

    err = ensure_built()
    if err:
        return {"impl": "reference", "unavailable": err}
    n = max(1, args.gpus)
    sys.path.insert(0, os.path.dirname(HERE))
    from swiftsnails_b200.models.word2vec import synthetic_corpus  # data generator only (shared corpus shape)

    sent_len = int(getattr(args, "sent_len", 1024))
    tokens = max(sent_len * 8, (args.ref_tokens // sent_len) * sent_len)
    wd = tempfile.mkdtemp(prefix="ssb200_refdata_")
    data_files = []
    for w in range(n):
        corpus = synthetic_corpus(args.vocab, tokens, s=getattr(args, "zipf_s", 1.0), seed=100 + w)
        path = os.path.join(wd, "data%d.txt" % w)
        with open(path, "w") as f:
            for s in range(tokens // sent_len):
                f.write(" ".join(map(str, corpus[s * sent_len:(s + 1) * sent_len])) + "\n")
        data_files.append(path)
    threads = _threads_per_worker(n)
    lr = float(getattr(args, "lr", 0.05))
    optimizer = getattr(args, "optimizer", "adagrad")
    neg_pool = int(getattr(args, "neg_pool", 16))
    prepop = int(getattr(args, "prepopulate", 1))
    sweep = [int(x) for x in str(getattr(args, "ref_minibatch_lines", "4,16")).split(",") if x]
    # a few hundred thousand tokens of a 1M-word vocabulary move the loss only a little (AdaGrad's first update of a
    # row is a full lr-sized step, which can lift the mean by a fraction of a percent): a run counts when its loss is
    # finite, stays at the untrained level ln2*(1+K) or below (2 % slack) and does not end above where it started
    untrained = math.log(2.0) * (1 + args.negatives)
    t0 = time.time()
    best, tried, last_err = None, [], None
    for mb in sweep:
        conf = {"dim": args.dim, "vocab": args.vocab, "window": args.window, "negative": args.negatives, "neg_pool": neg_pool,
                "minibatch_lines": mb, "learning_rate": lr, "optimizer": optimizer, "async_channel_thread_num": threads,
                "prepopulate": prepop}
        res, err = _one_run(n, conf, data_files)
        if err and threads > 1:
            # the reference's StateBarrier can lose a wake-up when many threads block at once: one retry, single-threaded
            conf["async_channel_thread_num"] = 1
            res, err = _one_run(n, conf, data_files)
        if err:
            last_err = err
            continue
        words = sum(r["words"] for r in res)
        t = max(r["train_s"] for r in res)
        rec = {"minibatch_lines": mb, "threads": int(res[0]["threads"]), "words_per_s": words / t,
               "mean_loss": sum(r["mean_loss"] for r in res) / len(res),
               "loss_first": sum(r["loss_first"] for r in res) / len(res), "loss_last": sum(r["loss_last"] for r in res) / len(res)}
        rec["trains"] = bool(math.isfinite(rec["mean_loss"]) and rec["mean_loss"] < 1.02 * untrained
                             and rec["loss_last"] <= 1.005 * rec["loss_first"])
        tried.append(rec)
        if rec["trains"] and (best is None or rec["words_per_s"] > best["words_per_s"]):
            best = rec
    wall = time.time() - t0
    shutil.rmtree(wd, ignore_errors=True)
    if best is None:
        why = last_err or ("reference arm does not train: " + json.dumps(tried))
        return {"impl": "reference", "unavailable": why[:400]}
    value = best["words_per_s"]
    return {
        "metric": "word2vec_sgns_words_per_sec", "value": value, "unit": "words/s", "n_gpus": n, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32 (CPU)", "data": "synthetic Zipf(s=1) corpus, random-init table", "impl": "reference",
        "config": {"model": "word2vec skip-gram negative sampling", "vocab": args.vocab, "dim": args.dim, "window": args.window,
                   "negatives": args.negatives, "neg_pool_per_128_pair_tile": neg_pool, "tokens_per_worker": tokens, "seq_len": sent_len,
                   "optimizer": optimizer, "lr": lr, "minibatch_lines": best["minibatch_lines"], "minibatch_sweep": tried,
                   "worker_threads": best["threads"], "data_path": "text corpus file read by scan_file_by_line inside the timed region",
                   "table": "all 2*vocab rows created before timing" if prepop else "empty at start, rows created lazily inside the timed region",
                   "parallelism": "1 master + %d servers + %d workers over ZeroMQ on the host CPUs" % (n, n),
                   "hash_map": "std::unordered_map stand-in for google::dense_hash_map",
                   "timing": "words / max over workers of the time inside train(); fixed sleeps excluded",
                   "mean_loss": best["mean_loss"], "loss_first_minibatch": best["loss_first"], "loss_last_minibatch": best["loss_last"],
                   "wall_s": wall},
        "e2e": {"value": value, "unit": "words/s", "note": "the reference is CPU-only: its timed region already reads the corpus file and pulls/pushes over TCP"},
        "gpu_launches": 0,
    }


if __name__ == "__main__":
    import argparse
    import json

    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0)
    ap.add_argument("--warmup", type=int, default=0)
    ap.add_argument("--vocab", type=int, default=1_000_000)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--window", type=int, default=5)
    ap.add_argument("--negatives", type=int, default=5)
    ap.add_argument("--ref-tokens", type=int, default=200_000)
    ap.add_argument("--ref-minibatch-lines", default="4,16")
    ap.add_argument("--lr", type=float, default=0.05)
    ap.add_argument("--optimizer", default="adagrad")
    ap.add_argument("--neg-pool", type=int, default=16)
    ap.add_argument("--prepopulate", type=int, default=1)
    ap.add_argument("--sent-len", type=int, default=1024)
    print(json.dumps(run(ap.parse_args())))
