#!/usr/bin/env python
"""Per-role view of an ncu report of the pipelined tensor-core SGNS kernel: warp-stall samples and executed
instructions aggregated by warp role (source-line ranges of csrc/kernels/sgns_tc.cu), the wait loops of every
mbarrier with the barrier they wait on, and the top source lines. Needs -lineinfo + --import-source on.
usage: python benchmarks/ncu_roles.py report.ncu-rep [n_tiles]"""
import collections, csv, io, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def page(rep, *extra):
    p = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + list(extra), capture_output=True, text=True)
    return list(csv.reader(io.StringIO(p.stdout)))


def main():
    rep = sys.argv[1]
    tiles = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    src = open(os.path.join(ROOT, "csrc", "kernels", "sgns_tc.cu")).read().split("\n")

    def find(s):
        for i, l in enumerate(src):
            if s in l:
                return i + 1
        return 10 ** 9
    marks = sorted([("producer", find("producer warp =====")), ("mma", find("MMA issue =====")), ("gather", find("cp.async gather warps (mode bit 32) =")),
                    ("e1", find("epilogue 1: logits")), ("e2", find("epilogue 2: gradients TMEM")), ("teardown", find("---- teardown"))], key=lambda kv: kv[1])

    def role(ln):
        r = "prologue"
        for name, l in marks:
            if ln >= l:
                r = name
        return r
    # SASS rows with their CUDA line: walk the cuda,sass view (a CUDA line row followed by its SASS rows)
    rows = page(rep, "--print-source", "cuda,sass")
    fname, hdr, cur = None, None, None
    by_role_s, by_role_i = collections.Counter(), collections.Counter()
    lines = {}
    waits = []
    for r in rows:
        if r and r[0] == "File Path":
            fname = os.path.basename(r[1]); continue
        if r and r[0] == "Line No":
            hdr = r; continue
        if not hdr or len(r) < 8:
            continue
        si, ii = hdr.index("# Samples"), hdr.index("Instructions Executed")
        if r[0].isdigit() and r[2] == "-":
            cur = (fname, int(r[0]), " ".join(r[1].split())[:100])
            n, ins = int(r[si] or 0), int(r[ii] or 0)
            e = lines.setdefault(cur[:2], [0, 0, cur[2]])
            e[0] += n; e[1] += ins
    # role attribution needs the inlining context: use the sass-only view, where rows are in address order, and the
    # cuda view's per-line totals for lines inside sgns_tc.cu; helper headers are attributed by address neighbourhood
    sass = page(rep)
    h = sass[1]
    ia, ii, isamp = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
    body = [r for r in sass[2:] if len(r) == len(h)]
    tot_s = sum(int(r[isamp] or 0) for r in body) or 1
    tot_i = sum(int(r[ii] or 0) for r in body) or 1
    print("kernel samples %d, instructions %d (%.0f per tile)" % (tot_s, tot_i, tot_i / tiles))
    print("\nwait loops (mbarrier address = smem offset; samples within +-6 instructions):")
    for k, r in enumerate(body):
        if "TRYWAIT" in r[ia] or "NANOSLEEP" in r[ia] or ("BAR.SYNC" in r[ia]):
            s = sum(int(x[isamp] or 0) for x in body[max(0, k - 2):k + 7])
            print("  %-74s exec/tile %8.1f  samples %5.1f%%" % (" ".join(r[ia].split())[:74], int(r[ii] or 0) / tiles, 100.0 * s / tot_s))
    print("\ntop CUDA lines by samples:")
    ts = sum(e[0] for e in lines.values()) or 1
    ti = sum(e[1] for e in lines.values()) or 1
    for (f, ln), (n, ins, s) in sorted(lines.items(), key=lambda kv: -kv[1][0])[:28]:
        print("  %5.1f%% smp %5.1f%% ins  %s:%d  %s" % (100.0 * n / ts, 100.0 * ins / ti, f, ln, s[:95]))
    for (f, ln), (n, ins, s) in lines.items():
        rl = role(ln) if f == "sgns_tc.cu" else f
        by_role_s[rl] += n; by_role_i[rl] += ins
    print("\nby role (inlined helper headers listed under their own file name):")
    for k, v in by_role_s.most_common():
        print("  %-26s samples %5.1f%%  instructions %5.1f%% (%d per tile)" % (k, 100.0 * v / ts, 100.0 * by_role_i[k] / ti, by_role_i[k] / tiles))


if __name__ == "__main__":
    main()
