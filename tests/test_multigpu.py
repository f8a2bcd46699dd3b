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
