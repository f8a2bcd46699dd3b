#!/usr/bin/env python
"""BASELINE config 3: sparse logistic regression, hashed 2^28 features, Criteo-shaped synthetic
batches (13 numeric + 26 categorical features per example), AdaGrad on the owner shard.
Prints examples/s and feature-keys/s (device-timed, max over ranks)."""
import argparse, json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.models.logistic_regression import LRConfig, LRWorker, criteo_like_batch
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.parallel.cluster import make_table

    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=262144)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--feature-bits", type=int, default=28)
    ap.add_argument("--fused", type=int, default=1)
    ap.add_argument("--feature-major", type=int, default=1, help="minibatches stored feature by feature ([39, B]): coalesced loads in the fused kernel")
    ap.add_argument("--optimizer", default="adagrad")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local); torch.cuda.set_device(dev)
    from swiftsnails_b200.parallel import group as ps_group
    grp = ps_group.init("master")            # ranks register with swift_master over ZeroMQ; no torch.distributed
    rank = grp.rank
    fm = bool(a.feature_major) and bool(a.fused)
    cfg = LRConfig(feature_bits=a.feature_bits, batch_size=a.batch, fused=bool(a.fused), optimizer=a.optimizer, feature_major=fm)
    # touched features are far fewer than 2^28; size the table for 64M rows in total
    table = make_table((1 << 26) // world, 1, dev, with_state=(2 if a.optimizer == "ftrl" else True), init_scale=0.0,
                       frag_num=max(1024, world))
    w = LRWorker(cfg, table, dev)
    n_batches = 4
    data = [criteo_like_batch(cfg, a.batch, seed=1000 * rank + b) for b in range(n_batches)]
    if fm:            # the data pipeline hands the minibatch over feature by feature
        data = [(np.ascontiguousarray(k.T), np.ascontiguousarray(v.T), y) for k, v, y in data]
    dev_data = [(u32_tensor(k, dev), torch.from_numpy(v).to(dev), torch.from_numpy(y).to(dev)) for k, v, y in data]
    for s in range(a.warmup):
        w.step(*dev_data[s % n_batches])
    torch.cuda.synchronize()
    grp.barrier()
    w.reset_loss()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(a.steps):
        w.step(*dev_data[s % n_batches])
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    ms = float(grp.all_reduce(float(ms), "max"))
    if rank == 0:
        ex = a.steps * a.batch * world
        print(json.dumps({"bench": "sparse_lr_adagrad", "n_gpus": world, "feature_bits": a.feature_bits, "batch_per_gpu": a.batch,
                          "features_per_example": cfg.n_feat, "ms_per_step": ms / a.steps, "examples_per_s": ex / (ms * 1e-3),
                          "feature_keys_per_s": ex * cfg.n_feat / (ms * 1e-3), "unique_keys_last_step": int(w.n_unique.item()),
                          "mean_loss": w.mean_loss(), "fused": bool(a.fused), "feature_major": fm, "optimizer": a.optimizer, "health": w.health(),
                          "phase_us_last_step": (w.fused.phase_us() if w.fused is not None else None)}))
    torch.cuda.synchronize()
    grp.barrier()
    grp.finish()
    grp.close()


if __name__ == "__main__":
    main()
