#!/usr/bin/env python
"""Kernel-structure probe for the inbox push on ONE GPU: R ranks emulated in one process (every shard and every inbox
segment is local memory), realistic sizes; times the ship kernel, the own-row update, and the segmented inbox apply
(+ mirror refresh) in isolation. What it cannot show is NVLink; what it shows is how far the kernels are from HBM speed."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.table import ShardedTable, round_capacity
    from swiftsnails_b200.parallel.cluster import Inbox

    R = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 304000
    pulls = (sys.argv[3] != "0") if len(sys.argv) > 3 else True
    dim, dev = 128, torch.device("cuda", 0)
    vocab = 2_000_000
    table = ShardedTable.local(round_capacity(vocab // R), dim, dev, num_shards=R, frag_num=1024, with_state=True, with_shadow=True)
    cap_rows = 2 * 1_100_000
    nbytes = Inbox.segment_bytes(cap_rows, dim, R, True)
    segs = [torch.zeros(nbytes, dtype=torch.uint8, device=dev) for _ in range(R)]
    ptrs = [t.data_ptr() for t in segs]
    inboxes = [Inbox(table, cap_rows, dev, bf16_rows=True, world=R, rank=r, segment_ptrs=ptrs, local_shard=table.owned[r],
                     owner_pulls=pulls) for r in range(R)]
    rng = np.random.default_rng(0)
    keysets = [u32_tensor(rng.choice(vocab, n, replace=False).astype(np.uint32), dev) for _ in range(R)]
    gslots = [table.pull(k, want_fp32=False)[1] for k in keysets]
    grads = [torch.randn(n, dim, device=dev) * 0.01 for _ in range(R)]
    for ib in inboxes:
        ib.own_stream = None
    out = {"ranks": R, "rows_per_rank": n, "owner_pulls": pulls}
    acc = {"send": 0.0, "own": 0.0, "handoff": 0.0, "apply": 0.0}
    reps = 8
    for it in range(reps + 2):
        ctxs = []
        for r in range(R):
            inboxes[r].marks = []
            ctxs.append(inboxes[r].ship(gslots[r], grads[r], 0.05, 1e-8, adagrad=True))
        for r in range(R):
            inboxes[r].apply(ctxs[r])
        torch.cuda.synchronize()
        if it >= 2:
            mk = inboxes[0].marks
            for (n0, e0), (n1, e1) in zip(mk[:-1], mk[1:]):
                if n1 in acc and not (n0 == "own" and n1 == "handoff" and R > 1 and False):
                    acc[n1] += e0.elapsed_time(e1) * 1e3 / reps
    # "handoff" of rank 0 spans the other ranks' ship halves in this emulation (one stream): not a kernel time
    out["us"] = {k: round(v, 1) for k, v in acc.items() if k != "handoff"}
    out["remote_rows_rank0"] = int((gslots[0].to(torch.int64) // table.capacity != 0).sum().item())
    out["bytes_shipped_rank0_MB"] = round(out["remote_rows_rank0"] * 256 / 1e6, 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
