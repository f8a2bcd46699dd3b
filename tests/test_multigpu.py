"""Multi-GPU tests (2 ranks, one per GPU): VMM peer-mapped shards, remote-load pull, remote-red
push, AdaGrad over NVLink, the p2p copy probe and a 2-GPU word2vec run.  Skipped with < 2 GPUs."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    try:
        import torch

        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_two_rank_peer_table(native_built):
    n = min(_ngpu(), int(os.environ.get("SSB200_TEST_GPUS", "2")))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "mgpu_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("MGPU_RESULT ")]
    assert line, p.stdout[-2000:]
    res = json.loads(line[0][len("MGPU_RESULT "):])
    assert res["pull_ok"] and res["push_ok"] and res["slot_cache_ok"] and res["p2p_read_gbs"] > 50
    assert res["inbox_ok"] and res["named_rule_ftrl_ok"] and res["fused_pull_ok"]
    print(res)


def _torchrun(n, port, args):
    return [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
            "127.0.0.1", "--master-port", str(port)] + args


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_two_rank_word2vec_cli_run_by_the_master(native_built, tmp_path):
    """register with swift_master over ZeroMQ -> train over NVLink -> WORKER_FINISH_WORK -> SERVER_TOLD_TO_TERMINATE ->
    one dump per rank; torch.distributed / NCCL is never initialised (ref: master/init.h, node_init.h, terminate.h)."""
    from swiftsnails_b200.parallel.launcher import parse_dump

    data, conf, out = str(tmp_path / "corpus.txt"), str(tmp_path / "worker.conf"), str(tmp_path / "params.txt")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_word2vec_data.py"), "--lines", "8000", "--vocab", "501",
                    "--zipf", "1.0", "--out", data], check=True)
    with open(conf, "w") as f:
        f.write("dim: 128\nvocab: 501\nwindow: 5\nnegative: 5\nlearning_rate: 0.05\nnum_iters: 2\nsent_len: 16\n"
                "tokens_per_step: 4096\nupdate: adagrad\nfrag_num: 64\n")
    env = dict(os.environ, SSB200_MASTER_LOG=str(tmp_path / "master.log"), SSB200_FORBID_TORCH_DIST="1")
    p = subprocess.run(_torchrun(2, 29541, ["-m", "swiftsnails_b200.train", "-config", conf, "-data", data, "-out", out]),
                       capture_output=True, text=True, timeout=900, cwd=ROOT, env=env)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    res = json.loads([l for l in p.stdout.splitlines() if l.startswith("RESULT ")][0][len("RESULT "):])
    assert res["gpus"] == 2 and res["control_plane"] == "master" and res["loss_per_pair"] < 4.0
    assert res["health_rank0"] == {"table_full": 0, "inbox_overflow": 0}
    keys = []
    for r in range(2):
        with open("%s.%d" % (out, r)) as f:
            d = parse_dump(f.read())
        assert len(d) > 100 and all(len(v) == 128 for v in d.values())
        keys += list(d)
    assert len(keys) == len(set(keys)) and 500 < len(keys) <= 2 * 501          # every key lives on exactly one shard
    log = open(str(tmp_path / "master.log")).read()
    assert "cluster is up (2 servers" in log and "Master terminated normally" in log


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_two_rank_lr_cli_run_by_the_master(native_built, tmp_path):
    import importlib.util

    spec = importlib.util.spec_from_file_location("_fw_tests", os.path.join(ROOT, "tests", "test_gpu_framework.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    _write_lr_examples = mod._write_lr_examples

    data, conf, out = str(tmp_path / "ex.txt"), str(tmp_path / "lr.conf"), str(tmp_path / "w.txt")
    _write_lr_examples(data, 30000, 12, 5000, 4)
    with open(conf, "w") as f:
        f.write("feature_bits: 16\nmax_features: 16\nminibatch: 3000\nupdate: ftrl\nlearning_rate: 0.2\nnum_iters: 4\n"
                "l1: 0.001\ntable_rows: 65536\nfrag_num: 64\n")
    p = subprocess.run(_torchrun(2, 29543, ["-m", "swiftsnails_b200.train", "-app", "lr", "-config", conf, "-data", data, "-out", out]),
                       capture_output=True, text=True, timeout=900, cwd=ROOT, env=dict(os.environ, SSB200_FORBID_TORCH_DIST="1"))
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    res = json.loads([l for l in p.stdout.splitlines() if l.startswith("RESULT ")][0][len("RESULT "):])
    assert res["gpus"] == 2 and res["control_plane"] == "master" and res["examples"] == 4 * 30000
    assert res["loss_last"] < 0.85 * res["loss_first"]
    seen = set()
    for line in open(data):
        seen |= set(int(t.split(":")[0]) for t in line.split()[1:])
    keys = []
    for r in range(2):
        keys += [int(l.split("\t")[0]) for l in open("%s.%d" % (out, r)).read().splitlines()]
    assert len(keys) == len(set(keys)) and set(keys) == seen
