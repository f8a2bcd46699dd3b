#!/usr/bin/env python
"""Measured NVLink bandwidth of kernel loads / stores on peer-mapped memory: the roofline denominators of the
multi-GPU pull (remote READ) and push (remote WRITE). One peer and all peers at once (NVSwitch: the per-GPU port
bandwidth is shared by however many peers are addressed), device-timed, max over ranks.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/nvlink_probe.py [--mb 256]
Prints one JSON line on rank 0; `probe()` is what benchmarks/pull_push_sweep.py calls for its denominators."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def probe(grp, dev, mb=256, iters=5):
    """{"read_one_gbs", "write_one_gbs", "read_all_gbs", "write_all_gbs"}: GB/s per GPU, max-time over ranks."""
    from swiftsnails_b200 import _lib
    from swiftsnails_b200.parallel.peer import PeerMemory

    world, rank = grp.world, grp.rank
    if world < 2:
        return None
    nbytes = (mb << 20) // 16 * 16
    pm = PeerMemory(nbytes, dev)
    local = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    streams = [torch.cuda.Stream(device=dev) for _ in range(world - 1)]
    peers = [(rank + 1 + i) % world for i in range(world - 1)]

    def run(targets, write):
        """One copy per target peer, all in flight together (one stream each); returns GB/s moved by this GPU."""
        per = nbytes // len(targets) // 16 * 16
        def once():
            cur = torch.cuda.current_stream()
            for st, p in zip(streams, targets):
                st.wait_stream(cur)
                with torch.cuda.stream(st):
                    off = targets.index(p) * per
                    src, dst = (local.data_ptr() + off, pm.segment_ptr(p) + off) if write else (pm.segment_ptr(p) + off, local.data_ptr() + off)
                    # every copy gets a full grid: the copies to different peers share the SMs, and a peer's share of
                    # the loads in flight must not shrink with the number of peers (a divided grid under-measures)
                    _lib.call("ssb_p2p_copy", src, dst, per, 148 * 4, _lib.stream_ptr())
            for st in streams[:len(targets)]:
                cur.wait_stream(st)
        for _ in range(2):
            once()
        torch.cuda.synchronize()
        grp.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            once()
        e1.record()
        torch.cuda.synchronize()
        ms = float(grp.all_reduce(e0.elapsed_time(e1) / iters, "max"))
        grp.barrier()
        return per * len(targets) / (ms * 1e-3) / 1e9

    out = {"n_gpus": world, "mb": mb,
           "read_one_gbs": run(peers[:1], False), "write_one_gbs": run(peers[:1], True),
           "read_all_gbs": run(peers, False), "write_all_gbs": run(peers, True),
           "nominal_gbs_per_direction": 900.0}
    del pm
    return out


def main():
    from swiftsnails_b200.parallel import group as ps_group

    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=256)
    a = ap.parse_args()
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    grp = ps_group.init("master")
    res = probe(grp, dev, a.mb)
    if grp.rank == 0:
        print(json.dumps(res))
    grp.barrier()
    grp.finish()
    grp.close()


if __name__ == "__main__":
    main()
