"""CPU tests of the C++ host core: unit-test binary, C-ABI bindings, and the automated localhost
cluster (1 master + 2 servers + 2 workers over ZeroMQ — BASELINE config 1)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "swiftsnails_b200", "bin")


def test_cpp_unit_tests(native_built):
    r = subprocess.run([os.path.join(BIN, "unit_tests")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failed expectations" in r.stdout


def test_config_python_matches_native(native_built, tmp_path):
    from swiftsnails_b200 import _lib
    from swiftsnails_b200.utils.config import ConfigParser

    (tmp_path / "common.conf").write_text("# c\nfrag_num: 100\nmaster_addr: tcp://127.0.0.1:16831\nfrag_num: 3\n")
    (tmp_path / "node.conf").write_text("import common.conf\nlisten_addr:\nshard_num: 8 # x\n")
    py = ConfigParser(str(tmp_path / "node.conf"))
    assert py.to_int32("frag_num") == 100 and py.get("master_addr") == "tcp://127.0.0.1:16831"
    assert py.get("listen_addr") == "" and py.to_int32("shard_num") == 8
    with pytest.raises(KeyError):
        py.get("nope")
    lib = _lib.core()
    lib.ssb_config_clear()
    assert lib.ssb_config_load(str(tmp_path / "node.conf").encode()) == 0
    buf = C.create_string_buffer(256)
    for k, v in py.items.items():
        n = lib.ssb_config_get(k.encode(), buf, 256)
        assert n >= 0 and buf.value.decode() == v
    assert lib.ssb_config_get(b"nope", buf, 256) == -1
    lib.ssb_config_clear()


def test_hashfrag_native_matches_python(native_built):
    from swiftsnails_b200 import _lib
    from swiftsnails_b200.parallel.hashfrag import HashFrag, fmix64

    lib = _lib.core()
    lib.ssb_hash_code.restype = C.c_uint64
    lib.ssb_hash_code.argtypes = [C.c_uint64]
    for x in (0, 1, 12345, 2 ** 40 + 7):
        assert lib.ssb_hash_code(x) == int(fmix64(np.array([x], dtype=np.uint64))[0])
    for nodes, frags in ((2, 100), (7, 100), (8, 1024)):
        hf = HashFrag(nodes, frags)
        m = np.zeros(frags, dtype=np.uint32)
        assert lib.ssb_hashfrag_build(nodes, frags, m.ctypes.data_as(C.c_void_p)) == 0
        assert np.array_equal(m, hf.map)
        keys = np.random.default_rng(0).integers(0, 2 ** 32, 5000, dtype=np.uint64).astype(np.uint32)
        out = np.zeros(len(keys), dtype=np.int32)
        lib.ssb_hashfrag_route(keys.ctypes.data_as(C.c_void_p), C.c_int64(len(keys)), m.ctypes.data_as(C.c_void_p),
                               frags, out.ctypes.data_as(C.c_void_p))
        assert np.array_equal(out, hf.to_node_id(keys))
    assert lib.ssb_hashfrag_build(8, 4, m.ctypes.data_as(C.c_void_p)) == -1
    assert lib.ssb_buffer_selftest() == 0


def test_text_dump_roundtrip(native_built, tmp_path):
    from swiftsnails_b200.utils import dump

    keys = np.array([5, 4220827, 4000000000], dtype=np.uint32)
    rows = np.array([[0.5, -1.25, 3.0], [0.012402, 1e-5, 7], [1, 2, 3]], dtype=np.float32)
    p = str(tmp_path / "param.txt")
    assert dump.write_text(p, keys, rows) == 3
    text = open(p).read().splitlines()
    assert text[0] == "5\tVec:\t0.5 -1.25 3 "          # the reference's Vec print form
    k2, r2 = dump.read_text(p, 3)
    assert np.array_equal(k2, keys) and np.allclose(r2, rows, rtol=1e-5)
    dump.write_text(p, keys[:1], rows[:1, :1].copy())
    assert open(p).read() == "5\t0.5\n"                  # scalar param: `key \t value`


@pytest.mark.parametrize("opt", ["sgd", "adagrad"])
def test_cpu_table_matches_reference_math(native_built, opt):
    from swiftsnails_b200.ops.cpu_table import CpuSparseTable
    from swiftsnails_b200.utils.reference import init_rows

    t = CpuSparseTable(16, shard_num=4, optimizer=opt, lr=0.1, eps=1e-8, seed=77, init_scale=0.5)
    keys = np.arange(100, 400, dtype=np.uint32)
    w = t.pull(keys)
    assert np.array_equal(w, init_rows(keys, 16, 77, 0.5))   # same lazy init as the HBM table
    acc = np.zeros_like(w)
    rng = np.random.default_rng(1)
    for _ in range(3):
        g = rng.standard_normal(w.shape).astype(np.float32)
        assert t.push(keys, g) == 0
        if opt == "sgd":
            w = w - 0.1 * g
        else:
            acc += g * g
            w = w - 0.1 * g / np.sqrt(acc + 1e-8)
    assert np.allclose(t.pull(keys), w, rtol=1e-5, atol=1e-6)
    assert t.push(np.array([999999], dtype=np.uint32), np.zeros((1, 16), np.float32)) == 1  # never pulled
    assert t.size() == 300


def test_localhost_cluster_1_master_2_servers_2_workers(native_built):
    """BASELINE config 1: synthetic sparse-key pull/push SGD over ZeroMQ, then check the final
    parameter dump against the closed form (init - lr * sum of pushed grads)."""
    from swiftsnails_b200.parallel.hashfrag import HashFrag
    from swiftsnails_b200.parallel.launcher import launch_local_cluster, parse_dump, parse_result_line
    from swiftsnails_b200.utils.reference import init_rows

    dim, iters, lr = 8, 3, 0.05
    r = launch_local_cluster("bench", 2, 2, overrides={"dim": dim, "bench_keys": 3000, "num_iters": iters,
                                                        "learning_rate": lr, "bench_key_space": 20000,
                                                        "frag_num": 100, "table_seed": 24301}, timeout=120)
    assert r["master"][0] == 0, r["master"][2]
    assert "Master terminated normally" in r["master"][2]
    for w in r["workers"]:
        assert w[0] == 0, w[2]
        res = parse_result_line(w[1], "BENCH")
        assert res and res["pull_keys_per_s"] > 0 and res["push_keys_per_s"] > 0
    dumps = [parse_dump(s[1]) for s in r["servers"]]
    assert all(s[0] == 0 for s in r["servers"])
    assert len(set(dumps[0]) & set(dumps[1])) == 0            # every key lives on exactly one server
    hf = HashFrag(2, 100)
    seen_ids = set()
    for d in dumps:
        ks = np.array(sorted(d), dtype=np.uint32)
        assert len(ks) > 500
        # same key -> server rule as the reference. Server ids are handed out in registration order
        # (whichever process reaches the master first is node 1), not launch order.
        ids = np.unique(hf.to_node_id(ks))
        assert len(ids) == 1 and int(ids[0]) in (1, 2)
        seen_ids.add(int(ids[0]))
        init = init_rows(ks, dim, 24301, 1.0 / dim)
        got = np.array([d[int(k)] for k in ks], dtype=np.float32)
        delta = (init - got) / (lr * 0.001)                   # = number of pushes that hit the key
        n_push = np.rint(delta[:, 0])
        assert np.all((n_push == iters) | (n_push == 2 * iters))  # pushed by one or by both workers
        assert np.allclose(delta, n_push[:, None], atol=0.02)
    assert seen_ids == {1, 2}


def test_cluster_word2vec_and_local_train(native_built, tmp_path):
    from swiftsnails_b200.parallel.launcher import launch_local_cluster, parse_result_line

    data = str(tmp_path / "data.txt")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_word2vec_data.py"), "--lines", "1500", "--vocab",
                    "301", "--out", data], check=True)
    ov = {"dim": 16, "vocab": 301, "window": 3, "negative": 3, "minibatch_lines": 32, "learning_rate": 0.05,
          "update": "adagrad", "num_iters": 2}
    r = launch_local_cluster("word2vec", 2, 2, overrides=ov, data_files=[data], timeout=240)
    assert r["master"][0] == 0, r["master"][2]
    for w in r["workers"]:
        assert w[0] == 0, w[2]
        res = parse_result_line(w[1])
        assert res["words"] > 10000 and res["words_per_s"] > 0 and 0 < res["mean_loss"] < 5.0
    # single-process mode (local_train=1): no servers, same algorithm
    conf = tmp_path / "local.conf"
    conf.write_text("".join("%s: %s\n" % kv for kv in dict(ov, local_train=1, num_iters=1).items()))
    p = subprocess.run([os.path.join(BIN, "swift_worker"), "-config", str(conf), "-data", data, "-app", "word2vec"],
                       capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    assert parse_result_line(p.stdout)["words"] > 10000


def test_lr_app_local(native_built, tmp_path):
    from swiftsnails_b200.parallel.launcher import parse_result_line

    rng = np.random.default_rng(0)
    w_true = rng.standard_normal(50)
    lines = []
    for _ in range(3000):
        f = rng.choice(50, 5, replace=False)
        y = 1 if w_true[f].sum() > 0 else 0
        lines.append("%d %s" % (y, " ".join("%d:1" % k for k in f)))
    data = tmp_path / "lr.txt"
    data.write_text("\n".join(lines) + "\n")
    conf = tmp_path / "lr.conf"
    conf.write_text("local_train: 1\nnum_iters: 5\nlearning_rate: 0.05\nminibatch_lines: 32\n")
    p = subprocess.run([os.path.join(BIN, "swift_worker"), "-config", str(conf), "-data", str(data), "-app", "lr"],
                       capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    assert parse_result_line(p.stdout)["mean_loss"] < 0.55   # below log(2): it learned


def test_corpus_loader(native_built, tmp_path):
    from swiftsnails_b200.utils.corpus import PAD, CorpusLoader

    lines = [" ".join(str(10 * i + j) for j in range(3 + i % 4)) for i in range(10)]
    p = tmp_path / "c.txt"
    p.write_text("\n".join(lines) + "\n")
    ld = CorpusLoader(str(p), batch_tokens=20, sent_len=5, epochs=2, pin=False)
    toks, real = [], 0
    for buf, n in ld:
        toks.append(buf.numpy().view(np.uint32).copy())
        real += n
    ld.close()
    allt = np.concatenate(toks)
    assert len(allt) == 20 * 5 and real == 2 * sum(min(len(l.split()), 5) for l in lines)
    sent0 = allt[:5]
    assert list(sent0[:3]) == [0, 1, 2] and all(sent0[3:] == PAD)          # padded to sent_len
    assert list(allt[5:10]) == [10, 11, 12, 13, PAD]
    assert list(allt[50:53]) == [0, 1, 2]                                   # second epoch starts over


def test_periodic_backup_files(native_built, tmp_path):
    """every `param_backup_period` push requests a server writes
    <param_backup_root>/param-<node id>-<counter>.txt in the dump format (one file name per server, D10)."""
    import glob

    from swiftsnails_b200.parallel.launcher import launch_local_cluster, parse_dump

    root = str(tmp_path / "backups")
    os.makedirs(root)
    r = launch_local_cluster("bench", 2, 1, overrides={"dim": 4, "bench_keys": 500, "num_iters": 4, "bench_key_space": 2000,
                                                       "frag_num": 50, "param_backup_period": 2, "param_backup_root": root},
                             timeout=120)
    assert r["master"][0] == 0, r["master"][2]
    files = sorted(glob.glob(os.path.join(root, "param-*-*.txt")))
    assert files, os.listdir(root)
    nodes = {os.path.basename(f).split("-")[1] for f in files}
    assert nodes == {"1", "2"}                                   # both servers, distinct file names
    final = [parse_dump(s[1]) for s in r["servers"]]
    for f in files:
        with open(f) as fh:
            d = parse_dump(fh.read())
        assert d and all(len(v) == 4 for v in d.values())
        assert any(set(d) <= set(fd) for fd in final)            # a backup is a snapshot of one server's keys


def test_master_start_up_timeout(native_built, tmp_path):
    """registration has a deadline. A master that never sees its nodes gives up after
    `master_time_out` seconds instead of waiting forever."""
    import time

    from swiftsnails_b200.parallel.launcher import BIN_DIR, write_configs

    wd = str(tmp_path)
    write_configs(wd, 1, 1, overrides={"master_time_out": 1})
    t0 = time.time()
    p = subprocess.run([os.path.join(BIN_DIR, "swift_master"), "-config", "master.conf"], cwd=wd, capture_output=True,
                       text=True, timeout=60)
    assert p.returncode != 0 and time.time() - t0 < 30
    assert "time" in (p.stderr + p.stdout).lower()


def test_dead_server_is_reported_not_hung(native_built, tmp_path):
    """SURVEY §5.3 (new framework): pull/push waits have a deadline (`rpc_timeout`), so a worker whose
    server died says so and exits instead of blocking forever in a barrier as the reference does."""
    import time

    from swiftsnails_b200.parallel.launcher import BIN_DIR, write_configs

    wd = str(tmp_path)
    write_configs(wd, 1, 1, overrides={"dim": 32, "bench_keys": 20000, "bench_key_space": 200000, "num_iters": 100000,
                                       "rpc_timeout": 2, "frag_num": 10})
    logs = {n: open(os.path.join(wd, n + ".log"), "w") for n in ("master", "server", "worker")}
    env = dict(os.environ, SSB200_LOCAL_IP="127.0.0.1")
    procs = {}
    try:
        procs["master"] = subprocess.Popen([os.path.join(BIN_DIR, "swift_master"), "-config", "master.conf"], cwd=wd,
                                           stdout=logs["master"], stderr=subprocess.STDOUT, env=env)
        time.sleep(0.2)
        procs["server"] = subprocess.Popen([os.path.join(BIN_DIR, "swift_server"), "-config", "node.conf"], cwd=wd,
                                           stdout=subprocess.DEVNULL, stderr=logs["server"], env=env)
        procs["worker"] = subprocess.Popen([os.path.join(BIN_DIR, "swift_worker"), "-config", "node.conf", "-app", "bench"],
                                           cwd=wd, stdout=logs["worker"], stderr=subprocess.STDOUT, env=env)
        time.sleep(1.5)                                  # registration done, the worker is pulling / pushing
        assert procs["worker"].poll() is None, "worker finished before the fault was injected"
        procs["server"].kill()
        rc = procs["worker"].wait(timeout=40)
        assert rc != 0
    finally:
        for p in procs.values():
            if p.poll() is None:
                p.kill()
            p.wait()
        for f in logs.values():
            f.close()
    with open(os.path.join(wd, "worker.log")) as f:
        out = f.read()
    assert "timed out" in out and "server is not answering" in out, out[-1500:]


def test_corpus_loader_split_and_cli_helpers(native_built, tmp_path):
    """Per-worker data split of the native loader (lines with index % n == i, empty lines not counted)
    and the host-side helpers of `python -m swiftsnails_b200.train` (config -> model config, step count)."""
    from swiftsnails_b200.train import count_sentences, parse_args, total_steps, word2vec_config
    from swiftsnails_b200.utils.config import ConfigParser
    from swiftsnails_b200.utils.corpus import PAD, CorpusLoader

    lines = []
    for i in range(23):
        lines.append(" ".join(str(100 * i + j) for j in range(1 + i % 3)))
        if i % 5 == 0:
            lines.append("")                                    # empty lines do not take an index
    p = tmp_path / "c.txt"
    p.write_text("\n".join(lines) + "\n")
    assert count_sentences(str(p)) == 23
    seen = []
    for rank in range(3):
        ld = CorpusLoader(str(p), batch_tokens=8, sent_len=4, epochs=1, pin=False, split=(3, rank))
        first = []
        for buf, n in ld:
            a = buf.numpy().view(np.uint32).reshape(-1, 4)
            first += [int(r[0]) // 100 for r in a if r[0] != PAD]
        ld.close()
        assert first == list(range(rank, 23, 3))                # worker `rank` gets sentences rank, rank+3, ...
        seen += first
    assert sorted(seen) == list(range(23))                      # disjoint and complete

    conf = tmp_path / "w.conf"
    conf.write_text("dim: 128\nvocab: 301\nwindow: 3\nnegative: 3\nlearning_rate: 0.1\nnum_iters: 2\n"
                    "sent_len: 4\ntokens_per_step: 10\nupdate: adagrad\n")
    cfg = word2vec_config(ConfigParser(str(conf)))
    assert (cfg.vocab, cfg.window, cfg.negatives, cfg.lr, cfg.optimizer) == (301, 3, 3, 0.1, "adagrad")
    assert cfg.tokens_per_step == 8 and cfg.tile_impl == "tcgen05"          # whole sentences per minibatch
    # 23 sentences, 3 workers: the smallest share is 7 per pass, 2 sentences per minibatch
    assert total_steps(23, 3, cfg, epochs=1) == 4 and total_steps(23, 3, cfg, epochs=2) == 7
    assert total_steps(2, 3, cfg) == 0
    # every worker really has that many minibatches
    for rank in range(3):
        ld = CorpusLoader(str(p), batch_tokens=8, sent_len=4, epochs=2, pin=False, split=(3, rank))
        assert sum(1 for _ in ld) >= total_steps(23, 3, cfg, epochs=2)
        ld.close()
    a = parse_args(["-config", "x.conf", "-data", "d.txt", "-out", "o.txt"])
    assert (a.config, a.data, a.out, a.app) == ("x.conf", "d.txt", "o.txt", "word2vec")
