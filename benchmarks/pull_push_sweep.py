#!/usr/bin/env python
"""Pull / push bandwidth sweep (BASELINE config 5): keys/s for 1K … 16M unique keys per batch,
ours (P2P kernels over NVLink peer memory) vs the NCCL all-to-all baseline, at the launched world
size.  Device-timed (CUDA events), max over ranks, >= 3 warm-up iterations per size.

  python benchmarks/pull_push_sweep.py [--dim 64] [--max-keys 16777216] [--optimizer sgd]
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/pull_push_sweep.py
One JSON line per (impl, size) on rank 0; roofline fractions are computed against the measured
HBM copy bandwidth (MEASURED_PEAKS.json) and, with several GPUs, against NVLink bandwidths MEASURED at the start of
the run by benchmarks/nvlink_probe.py (all-peers kernel reads for the pull, all-peers kernel writes for the push);
the fraction against the nominal 900 GB/s per direction is printed next to it. The NCCL arm (`nccl`: split sizes
through the host; `nccl_padded`: fixed-capacity buckets, no host sync, bf16 rows on the wire) runs in a
torch.distributed job, each phase timed with its own CUDA events.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.models.embedding_bench import PullPushWorker
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.parallel.cluster import make_table
    from swiftsnails_b200.parallel.nccl_baseline import NcclShardedTable

    ap = argparse.ArgumentParser()
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--min-keys", type=int, default=1024)
    ap.add_argument("--max-keys", type=int, default=16 * 1024 * 1024)
    ap.add_argument("--key-space", type=int, default=0, help="distinct keys overall (default 2 x max-keys x world)")
    ap.add_argument("--optimizer", default="sgd", choices=["sgd", "adagrad"])
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--impls", default="p2p,nccl,nccl_padded")
    ap.add_argument("--out", default="")
    a = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    from swiftsnails_b200.parallel import group as ps_group
    grp = ps_group.init("torch")           # the NCCL arms need torch.distributed; the p2p arm only uses the group for set-up
    key_space = a.key_space or 2 * a.max_keys * world
    table = make_table((key_space + world - 1) // world, a.dim, dev, with_state=(a.optimizer == "adagrad"),
                       init_scale=0.1, frag_num=max(1024, world))
    tables = {"p2p": table}
    if "nccl" in a.impls.split(","):
        tables["nccl"] = NcclShardedTable(table)
    if "nccl_padded" in a.impls.split(","):
        tables["nccl_padded"] = NcclShardedTable(table, padded=True, wire_bf16=True)
    hbm = 6541.8e9
    try:
        hbm = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"] * 1e9
    except Exception:
        pass
    nv = None
    if world > 1:
        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        from nvlink_probe import probe
        nv = probe(grp, dev)
        if rank == 0:
            print(json.dumps({"nvlink_probe": nv}), flush=True)
    # the better of the two probe shapes: with many peers the probe's per-peer copy kernels crowd one another and
    # under-measure (8 GPUs: 349 GB/s "all peers" vs 624 GB/s one peer, while the pull kernel itself sustains ~620)
    nv_read = (max(nv["read_all_gbs"], nv["read_one_gbs"]) if nv else 900.0) * 1e9
    nv_write = (max(nv["write_all_gbs"], nv["write_one_gbs"]) if nv else 900.0) * 1e9

    def timed(fn, iters):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    results = []
    n = a.min_keys
    rng = np.random.default_rng(1000 + rank)
    while n <= a.max_keys:
        keys_np = rng.choice(key_space, size=n, replace=False).astype(np.uint32) if n * 4 <= key_space else \
            np.unique(rng.integers(0, key_space, size=int(n * 1.2), dtype=np.int64))[:n].astype(np.uint32)
        n_eff = len(keys_np)
        keys = u32_tensor(keys_np, dev)
        for impl in a.impls.split(","):
            t = tables[impl]
            w = PullPushWorker(t, n_eff, dev, optimizer=a.optimizer)
            for _ in range(3):
                w.pull(keys)
                w.push(n_eff)
            ms_pull = timed(lambda: w.pull(keys), a.iters)
            ms_push = timed(lambda: w.push(n_eff), a.iters)   # NCCL: re-uses the exchange plan of the last pull
            row_b = a.dim * 4
            remote = (world - 1) / world
            # bytes that must move per key: pull = key-bucket probe (128 B) + row read + row write;
            # push = grad read + row read-modify-write (+ state RMW for AdaGrad)
            pull_bytes = n_eff * (128 + 2 * row_b)
            push_bytes = n_eff * (row_b * (3 + (2 if a.optimizer == "adagrad" else 0)))
            roof_pull = max(pull_bytes / hbm, n_eff * remote * (128 + row_b) / nv_read) if world > 1 else pull_bytes / hbm
            roof_push = max(push_bytes / hbm, n_eff * remote * row_b / nv_write) if world > 1 else push_bytes / hbm
            nom_pull = max(pull_bytes / hbm, n_eff * remote * (128 + row_b) / 900e9) if world > 1 else pull_bytes / hbm
            nom_push = max(push_bytes / hbm, n_eff * remote * row_b / 900e9) if world > 1 else push_bytes / hbm
            res = {"impl": impl, "n_gpus": world, "keys_per_batch_per_gpu": n_eff, "dim": a.dim, "optimizer": a.optimizer,
                   "pull_ms": ms_pull, "push_ms": ms_push,
                   "pull_keys_per_s_total": n_eff * world / (ms_pull * 1e-3),
                   "push_keys_per_s_total": n_eff * world / (ms_push * 1e-3),
                   "pull_roofline_frac": roof_pull / (ms_pull * 1e-3), "push_roofline_frac": roof_push / (ms_push * 1e-3),
                   "pull_frac_of_nominal_900": nom_pull / (ms_pull * 1e-3), "push_frac_of_nominal_900": nom_push / (ms_push * 1e-3),
                   "nvlink_read_gbs_measured": nv_read / 1e9 if nv else None,
                   "nvlink_write_gbs_measured": nv_write / 1e9 if nv else None}
            results.append(res)
            if rank == 0:
                print(json.dumps(res), flush=True)
            del w
        n *= 4
    if rank == 0 and a.out:
        with open(a.out, "w") as f:
            for r in results:
                f.write(json.dumps(r) + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
