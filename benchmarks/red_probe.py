#!/usr/bin/env python
"""Reduction-throughput probe: fp32 row adds into a buffer shaped like the worker's gradient cache
(304 K rows x 512 B by default), issued from every SM at random rows. Tells how many bytes per second
of `red` the L2 atomic units take for each instruction form — the floor under the SGNS kernel's
epilogue 2 (csrc/kernels/p2p.cu: red_probe_kernel)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from swiftsnails_b200 import _lib
    dev = torch.device("cuda", 0)
    n_rows = int(sys.argv[1]) if len(sys.argv) > 1 else 304_000
    iters = 128
    buf = torch.zeros((n_rows, 128), dtype=torch.float32, device=dev)
    names = {0: "red.v4.f32, 512 B row per warp", 1: "4 x scalar red.f32, 128 B each", 2: "red.v4.f32, 4 rows x 128 B",
             3: "plain v4 stores, 512 B row per warp"}
    out = {"n_rows": n_rows, "buffer_MB": n_rows * 512 / 1e6}
    for blocks in (148, 296):
        for pat in (0, 1, 2, 3):
            ts = []
            for rep in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                _lib.call("ssb_red_probe", buf.data_ptr(), n_rows, iters, pat, blocks, _lib.stream_ptr())
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ms = sorted(ts)[len(ts) // 2]
            nbytes = blocks * 16 * iters * 512
            out["blocks%d_pattern%d" % (blocks, pat)] = {"what": names[pat], "ms": round(ms, 4), "GB_per_s": round(nbytes / ms / 1e6, 1),
                                                         "rows_per_us": round(blocks * 16 * iters / ms / 1e3, 1)}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
