#!/usr/bin/env python
"""Pull/push keys/s of the UNMODIFIED reference (ZeroMQ, CPUs) and of this repo's own CPU/ZeroMQ
path for a sweep of keys per batch: 1 master + N servers + N workers on localhost (config 5's
"vs ref ZeroMQ" column).  usage: ref_pull_push_sweep.py [--nodes 2] [--dim 128] [--sizes 1024,16384,262144]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from baseline import run_reference as rr
    from swiftsnails_b200.parallel.launcher import launch_local_cluster, parse_result_line

    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=2)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--sizes", default="1024,16384,262144")
    ap.add_argument("--iters", type=int, default=3)
    a = ap.parse_args()
    err = rr.ensure_built()
    for n_keys in [int(x) for x in a.sizes.split(",")]:
        conf = {"dim": a.dim, "bench_keys": n_keys, "bench_key_space": max(4 * n_keys * a.nodes, 1 << 20), "num_iters": a.iters,
                "vocab": 1, "window": 1, "negative": 1, "minibatch_lines": 1, "learning_rate": 0.05}
        row = {"keys_per_batch_per_worker": n_keys, "dim": a.dim, "servers": a.nodes, "workers": a.nodes}
        if err is None:
            outs, failed, _ = rr.launch(a.nodes, "bench", conf, timeout=3600)
            if not failed:
                res = [rr._parse(o, "BENCH") for o in outs]
                row["reference_pull_keys_per_s_total"] = sum(r["pull_keys_per_s"] for r in res)
                row["reference_push_keys_per_s_total"] = sum(r["push_keys_per_s"] for r in res)
            else:
                row["reference_error"] = failed[:200]
        r = launch_local_cluster("bench", a.nodes, a.nodes, overrides={k: v for k, v in conf.items() if k in ("dim", "bench_keys", "bench_key_space", "num_iters", "learning_rate")}, timeout=3600)
        res = [parse_result_line(w[1], "BENCH") for w in r["workers"]]
        if all(res):
            row["ours_cpu_pull_keys_per_s_total"] = sum(x["pull_keys_per_s"] for x in res)
            row["ours_cpu_push_keys_per_s_total"] = sum(x["push_keys_per_s"] for x in res)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
