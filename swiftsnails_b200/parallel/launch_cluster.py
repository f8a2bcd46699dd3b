"""CLI: python -m swiftsnails_b200.parallel.launch_cluster --app bench --servers 2 --workers 2"""
import argparse
import sys

from .launcher import launch_local_cluster


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--app", default="bench")
    ap.add_argument("--servers", type=int, default=2)
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--data", nargs="*", default=None)
    ap.add_argument("--set", nargs="*", default=[], help="config overrides key=value")
    a = ap.parse_args()
    ov = dict(kv.split("=", 1) for kv in a.set)
    r = launch_local_cluster(a.app, a.servers, a.workers, overrides=ov, data_files=a.data)
    print("master rc", r["master"][0], "| %.2f s | workdir %s" % (r["seconds"], r["workdir"]))
    for i, w in enumerate(r["workers"]):
        print("worker %d rc %d: %s" % (i, w[0], w[1].strip().splitlines()[-1] if w[1].strip() else ""))
    ok = r["master"][0] == 0 and all(w[0] == 0 for w in r["workers"]) and all(s[0] == 0 for s in r["servers"])
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
