#!/usr/bin/env python
"""Per-stage device time of the unfused sparse-LR step (dedup -> pull -> fwd/bwd -> push) on one GPU."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.models.logistic_regression import LRConfig, LRWorker, criteo_like_batch
    from swiftsnails_b200.ops import lr as lr_ops
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.parallel.cluster import make_table

    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
    dev = torch.device("cuda", 0)
    cfg = LRConfig(batch_size=batch)
    table = make_table(1 << 26, 1, dev, with_state=True, init_scale=0.0)
    w = LRWorker(cfg, table, dev)
    data = [criteo_like_batch(cfg, batch, seed=b) for b in range(3)]
    dd = [(u32_tensor(k, dev), torch.from_numpy(v).to(dev), torch.from_numpy(y).to(dev)) for k, v, y in data]
    for s in range(6):
        w.step(*dd[s % 3])
    torch.cuda.synchronize()
    names = ["dedup", "pull", "fwd_bwd", "push"]
    acc = {n: 0.0 for n in names}
    reps = 12
    for s in range(reps):
        k, v, y = dd[s % 3]
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        ev[0].record()
        w.dedup(k.reshape(-1), unique=w.unique, inverse=w.inverse, n_unique=w.n_unique)
        ev[1].record()
        w.table.pull(w.unique, n_dev=w.n_unique, rows_out=w.params, gslot_out=w.gslot, grads_zero=w.grads, n=w.n)
        ev[2].record()
        lr_ops.lr_fwd_bwd(w.params, w.grads, w.inverse, v, y, cfg.n_feat, w.loss)
        ev[3].record()
        w.table.push_adagrad(w.gslot, w.grads, cfg.lr, cfg.adagrad_eps, n_dev=w.n_unique, zero_grads=False, n=w.n)
        ev[4].record()
        torch.cuda.synchronize()
        for i, n in enumerate(names):
            acc[n] += ev[i].elapsed_time(ev[i + 1]) / reps
    print(json.dumps({"batch": batch, "unique": int(w.n_unique.item()), "instances": batch * cfg.n_feat,
                      "ms": {k: round(v, 4) for k, v in acc.items()}, "total_ms": round(sum(acc.values()), 4)}))


if __name__ == "__main__":
    main()
