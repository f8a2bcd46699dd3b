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
