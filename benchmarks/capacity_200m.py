#!/usr/bin/env python
"""BASELINE config 4: large embedding table, 200M rows x 64 fp32 with AdaGrad state, sharded across
the GPUs (25M rows/GPU at 8 GPUs = 12.9 GB rows+state+keys per GPU at load factor 0.5 -> 25.8 GB of
slots).  Each step every worker pulls and AdaGrad-pushes a batch of random keys drawn from the
whole key space; rows are created lazily on first touch.  `--rows-total` scales it down for fewer
GPUs (default: 25M rows per launched GPU, i.e. the per-GPU share of the 200M-row config)."""
import argparse, json, os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.models.embedding_bench import PullPushWorker
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.parallel.cluster import make_table

    ap = argparse.ArgumentParser()
    ap.add_argument("--rows-total", type=int, default=0)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--batch", type=int, default=1 << 20)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local); torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rows_total = a.rows_total or 25_000_000 * world
    table = make_table((rows_total + world - 1) // world, a.dim, dev, with_state=True, init_scale=0.01, frag_num=max(1024, world))
    shard_gb = table.owned[0].nbytes() / 2 ** 30
    w = PullPushWorker(table, a.batch, dev, optimizer="adagrad", lr=0.01)
    g = torch.Generator(device=dev); g.manual_seed(1234 + rank)
    batches = [torch.randint(0, rows_total, (a.batch,), generator=g, device=dev, dtype=torch.int64).to(torch.int32) for _ in range(8)]
    for s in range(a.warmup):
        w.dedup_pull_push(batches[s % 8])
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(a.steps):
        w.dedup_pull_push(batches[s % 8])
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    live = torch.tensor([table.size()], device=dev, dtype=torch.int64)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
        dist.all_reduce(live)
    if rank == 0:
        keys = a.steps * a.batch * world
        print(json.dumps({"bench": "embedding_capacity_adagrad", "n_gpus": world, "rows_total_capacity": rows_total, "dim": a.dim,
                          "shard_gib": shard_gb, "batch_per_gpu": a.batch, "ms_per_step": ms / a.steps,
                          "keys_per_s_pull_plus_push": keys / (ms * 1e-3), "live_rows": int(live.item()),
                          "mem_allocated_gib": torch.cuda.max_memory_allocated(dev) / 2 ** 30}))
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
