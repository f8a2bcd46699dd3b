"""The reference-shaped user API on the GPU path: BaseAlgorithm / SwiftWorker / SwiftServer,
DeviceParamCache, pull_with_barrier / push_with_barrier, access methods, backup + dump + reload."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_user_algorithm_roundtrip(device, tmp_path):
    from swiftsnails_b200.framework import (AdaGradPushMethod, BaseAlgorithm, DeviceParamCache, SgdPushMethod,
                                            SwiftServer, SwiftWorker, pull_with_barrier, push_with_barrier)
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable
    from swiftsnails_b200.utils import dump
    from swiftsnails_b200.utils.config import ConfigParser
    from swiftsnails_b200.utils.reference import init_rows

    (tmp_path / "worker.conf").write_text("num_iters: 3\nlearning_rate: 0.1\nparam_backup_period: 2\nparam_backup_root: %s\n" % tmp_path)
    cfg = ConfigParser(str(tmp_path / "worker.conf"))
    table = ShardedTable.local(4096, 8, device, num_shards=2, frag_num=16, with_state=True, seed=5, init_scale=0.5)
    server = SwiftServer(table, cfg)

    class Alg(BaseAlgorithm):
        """Pull a key set (with duplicates), add a constant gradient per occurrence, push with SGD."""
        def train(self):
            self.keys = np.array([3, 7, 7, 11, 3, 3, 99], dtype=np.uint32)
            cache = DeviceParamCache(64, 8, device)
            method = SgdPushMethod(self.config.to_float("learning_rate"))
            for _ in range(self.config.to_int32("num_iters")):
                pull_with_barrier(self.table, u32_tensor(self.keys, device), cache)
                ones = torch.ones(len(self.keys), 8, device=device)
                cache.grads.index_add_(0, cache.inverse[:len(self.keys)].long(), ones)   # accumulate per unique key
                push_with_barrier(self.table, cache, method)
                server.on_push()
            self.cache = cache

    alg = Alg()
    SwiftWorker(alg, table, cfg)()
    uniq, counts = np.unique(alg.keys, return_counts=True)
    rows, _ = table.pull(u32_tensor(uniq, device), insert=False)
    expect = init_rows(uniq, 8, 5, 0.5) - 0.1 * 3 * counts[:, None].astype(np.float32)
    assert np.allclose(rows.cpu().numpy(), expect, atol=1e-5)
    assert float(alg.cache.grads.abs().max()) == 0.0                       # Grad::reset() after push
    # periodic backup every 2 pushes, in the reference's text format
    backup = tmp_path / "param-0-2.txt"
    assert backup.exists()
    assert all("\tVec:\t" in l for l in backup.read_text().splitlines())
    # final dump -> reload into a fresh table (checkpoint resume)
    final = str(tmp_path / "final.txt")
    assert server.terminate(final) == len(uniq)
    t2 = ShardedTable.local(4096, 8, device, num_shards=3, frag_num=9)
    assert dump.load_table(t2, final) == len(uniq)
    rows2, _ = t2.pull(u32_tensor(uniq, device), insert=False)
    assert np.allclose(rows2.cpu().numpy(), expect, rtol=1e-4, atol=1e-5)   # %g text precision
    # AdaGrad method through the same API
    cache = DeviceParamCache(64, 8, device)
    pull_with_barrier(table, u32_tensor(uniq, device), cache)
    cache.grads[:len(uniq)] = 2.0
    push_with_barrier(table, cache, AdaGradPushMethod(0.1, 1e-8))
    rows3, _ = table.pull(u32_tensor(uniq, device), insert=False)
    assert np.allclose(rows3.cpu().numpy(), expect - 0.1 * 2.0 / np.sqrt(4.0 + 1e-8), atol=1e-5)
    # binary dump alternative
    binp = str(tmp_path / "shard.bin")
    assert dump.save_table_binary(table, binp) > 0
    t3 = ShardedTable.local(4096, 8, device)
    assert dump.load_table_binary(t3, binp) == len(uniq)


def test_train_from_text_corpus_with_ragged_sentences(device, tmp_path):
    """Native corpus loader -> pinned batches -> tcgen05 step; padding tokens must be ignored."""
    import subprocess
    import sys

    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecTrainer
    from swiftsnails_b200.ops.table import ShardedTable, round_capacity

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    data = str(tmp_path / "corpus.txt")
    subprocess.run([sys.executable, os.path.join(root, "tools", "gen_word2vec_data.py"), "--lines", "6000", "--vocab", "301",
                    "--zipf", "1.0", "--out", data], check=True)                      # 6-15 ids per line, like the reference's generator
    cfg = Word2VecConfig(vocab=301, dim=128, window=5, negatives=5, neg_pool=16, sent_len=16, tokens_per_step=4096,
                         lr=0.05, optimizer="adagrad", tile_impl="tcgen05")
    table = ShardedTable.local(round_capacity(2 * cfg.vocab), cfg.dim, device, with_state=True, init_scale=1.0 / cfg.dim)
    tr = Word2VecTrainer(cfg, table, device)
    n_lines = sum(1 for _ in open(data))
    n_tokens = sum(len(l.split()) for l in open(data))
    first = None
    total, last = tr.train_file(data, epochs=3)
    assert total == 3 * n_tokens                                                     # every real token, no padding counted
    assert table.size() <= 2 * cfg.vocab                                             # no row was created for the padding key
    assert np.isfinite(last) and last < 4.0, last                                    # 6*ln2 = 4.16 at init


def test_train_cli_one_gpu(device, tmp_path):
    """`python -m swiftsnails_b200.train -config … -data … -out …`: the GPU counterpart of
    `swift_worker -config … -data …` — trains from a text corpus and leaves a reference-format dump."""
    import json
    import subprocess
    import sys

    from swiftsnails_b200.parallel.launcher import parse_dump

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    data, conf, out = str(tmp_path / "corpus.txt"), str(tmp_path / "worker.conf"), str(tmp_path / "params.txt")
    subprocess.run([sys.executable, os.path.join(root, "tools", "gen_word2vec_data.py"), "--lines", "6000", "--vocab", "301",
                    "--zipf", "1.0", "--out", data], check=True)
    with open(conf, "w") as f:
        f.write("dim: 128\nvocab: 301\nwindow: 5\nnegative: 5\nlearning_rate: 0.05\nnum_iters: 3\nsent_len: 16\n"
                "tokens_per_step: 4096\nupdate: adagrad\nparam_backup_period: 4\nparam_backup_root: %s\n" % str(tmp_path / "backups"))
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    p = subprocess.run([sys.executable, "-m", "swiftsnails_b200.train", "-config", conf, "-data", data, "-out", out],
                       capture_output=True, text=True, timeout=600, cwd=root, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    res = json.loads([l for l in p.stdout.splitlines() if l.startswith("RESULT ")][0][len("RESULT "):])
    n_tokens = sum(len(l.split()) for l in open(data))
    assert res["tokens"] == 3 * n_tokens and res["loss_per_pair"] < 4.0 and res["words_per_s"] > 0
    with open(out) as f:
        d = parse_dump(f.read())
    assert 300 < len(d) <= 2 * 301 and all(len(v) == 128 for v in d.values())
    assert all(0 <= k < 2 * 301 for k in d)
    # periodic backups (reference: every param_backup_period push requests, system/server/init.h:128-149)
    backups = sorted(os.listdir(str(tmp_path / "backups")))
    assert res["backups_rank0"] == len(backups) >= 1 and all(b.startswith("param-0-") for b in backups)
    assert res["health_rank0"] == {"table_full": 0, "inbox_overflow": 0} and res["ids_out_of_vocab_rank0"] == 0
    # resume from a dump that carries the optimizer state, train one more pass
    ck = str(tmp_path / "ck.txt")
    p = subprocess.run([sys.executable, "-m", "swiftsnails_b200.train", "-config", conf, "-data", data, "-out", ck, "--dump-state",
                        "--epochs", "1"], capture_output=True, text=True, timeout=600, cwd=root, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "\tAcc:\t" in open(ck).readline()
    p = subprocess.run([sys.executable, "-m", "swiftsnails_b200.train", "-config", conf, "-data", data, "-out", out, "--resume", ck,
                        "--epochs", "1"], capture_output=True, text=True, timeout=600, cwd=root, env=env)
    assert p.returncode == 0 and "resumed" in p.stdout, p.stdout[-2000:] + p.stderr[-2000:]


def _write_lr_examples(path, n, n_feat, space, seed):
    """`label key:value …` lines from a hidden sparse linear model (so there is something to learn)."""
    rng = np.random.default_rng(seed)
    w = rng.standard_normal(space) * 1.5
    with open(path, "w") as f:
        for _ in range(n):
            ks = np.unique((space * rng.random(n_feat) ** 2).astype(np.int64))
            vs = np.where(rng.random(len(ks)) < 0.3, np.round(rng.random(len(ks)) * 2, 3), 1.0)
            y = int(rng.random() < 1.0 / (1.0 + np.exp(-float((w[ks] * vs).sum()))))
            f.write(str(y) + " " + " ".join(("%d" % k) if v == 1.0 else ("%d:%g" % (k, v)) for k, v in zip(ks, vs)) + "\n")


@pytest.mark.parametrize("update", ["adagrad", "ftrl"])
def test_train_cli_lr_app_one_gpu(device, tmp_path, update):
    """`-app lr`: sparse logistic regression from a `label key:value …` file through the fused GPU step; the dump
    has one weight per seen feature, the loss falls."""
    import json
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    data, conf, out = str(tmp_path / "ex.txt"), str(tmp_path / "lr.conf"), str(tmp_path / "w.txt")
    _write_lr_examples(data, 20000, 12, 5000, 3)
    with open(conf, "w") as f:
        f.write("feature_bits: 16\nmax_features: 16\nminibatch: 2000\nupdate: %s\nlearning_rate: %s\nnum_iters: 4\n"
                "l1: 0.001\ntable_rows: 65536\n" % (update, "0.1" if update == "adagrad" else "0.2"))
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    p = subprocess.run([sys.executable, "-m", "swiftsnails_b200.train", "-app", "lr", "-config", conf, "-data", data, "-out", out],
                       capture_output=True, text=True, timeout=600, cwd=root, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    res = json.loads([l for l in p.stdout.splitlines() if l.startswith("RESULT ")][0][len("RESULT "):])
    assert res["app"] == "lr" and res["examples"] == 4 * 20000 and res["update"] == update
    assert res["loss_last"] < 0.8 * res["loss_first"] and res["loss_first"] < 0.75
    assert res["health_rank0"]["table_full"] == 0 and res["health_rank0"]["touched_overflow"] == 0
    seen = set()
    for line in open(data):
        seen |= set(int(t.split(":")[0]) for t in line.split()[1:])
    lines = open(out).read().splitlines()
    assert len(lines) == len(seen) == res["rows_dumped_rank0"]
    keys = set(int(l.split("\t")[0]) for l in lines)
    assert keys == seen


@pytest.mark.parametrize("impl", ["cuda_core", "tcgen05"])
def test_word2vec_learns_planted_structure(device, impl):
    """Quality check, not just 'loss goes down': in a corpus where word a is always followed by
    partner(a), the trained score in[a].out[partner(a)] must beat in[a].out[random] for nearly
    every a — for both the per-pair CUDA-core kernel and the tensor-core block formulation."""
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable, round_capacity

    V, B = 1000, 8192
    cfg = Word2VecConfig(vocab=V, dim=128, window=1, negatives=5, neg_pool=16, sent_len=64, tokens_per_step=B,
                         lr=0.1, optimizer="adagrad", tile_impl=impl)
    table = ShardedTable.local(round_capacity(2 * V), cfg.dim, device, with_state=True, init_scale=1.0 / cfg.dim)
    w = Word2VecWorker(cfg, table, device)
    rng = np.random.default_rng(0)
    partner = rng.permutation(V)
    for step in range(100):
        a = rng.integers(0, V, B // 2)
        toks = np.empty(B, dtype=np.uint32)
        toks[0::2], toks[1::2] = a, partner[a]
        w.step(u32_tensor(toks, device))
    torch.cuda.synchronize()
    words = np.arange(V, dtype=np.uint32)
    vin, _ = table.pull(u32_tensor(words, device), insert=False)
    vout, _ = table.pull(u32_tensor(words + V, device), insert=False)
    scores = vin @ vout.t()                                        # [V, V]
    pos = scores[torch.arange(V), torch.from_numpy(partner).to(device)]
    rnd = scores[torch.arange(V), torch.from_numpy(rng.permutation(V)).to(device)]
    frac = float((pos > rnd).float().mean())
    top1 = float((scores.argmax(1).cpu() == torch.from_numpy(partner)).float().mean())
    print(impl, "pos>random", frac, "top-1", top1)
    assert frac > 0.97 and top1 > 0.3, (impl, frac, top1)
