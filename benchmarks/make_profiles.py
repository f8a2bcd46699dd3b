#!/usr/bin/env python
"""Turn the ncu reports brought back in gpurun_out/ into the tracked summaries under profiles/:
selected raw metrics per kernel, the top stall lines of the source page, SASS mnemonic evidence
(UTC*MMA / LDTM / UTMALDG / REDG …) and the full SASS of the tensor-core kernel.
usage: python benchmarks/make_profiles.py [report.ncu-rep …]"""
import collections
import csv
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
LIB = os.path.join(ROOT, "swiftsnails_b200", "lib", "libssb200_kernels.so")
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum",
        "lts__t_sector_hit_rate.pct", "launch__occupancy_limit_registers", "sm__maximum_warps_per_active_cycle_pct"]


def ncu_csv(rep, page):
    p = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True)
    return list(csv.reader(io.StringIO(p.stdout)))


def raw_summary(rep):
    rows = ncu_csv(rep, "raw")
    if len(rows) < 3:
        return []
    hdr, units, body = rows[0], rows[1], rows[2:]
    name_i = hdr.index("Kernel Name")
    out = []
    for r in body:
        d = {"kernel": r[name_i]}
        for i, h in enumerate(hdr):
            if h in WANT:
                d[h] = "%s %s" % (r[i], units[i])
        out.append(d)
    return out


def stalls(rep, top=25):
    rows = ncu_csv(rep, "source")
    res, cur, hdr = [], None, None
    for r in rows:
        if "Source" in r and "Address" in r:
            hdr = {h: i for i, h in enumerate(r)}
            cur = []
            res.append(cur)
            continue
        if hdr is None or len(r) < len(hdr):
            continue
        try:
            v = float(r[hdr["Warp Stall Sampling (All Samples)"]])
        except ValueError:
            continue
        cur.append((v, r[hdr["Source"]], r[hdr["Instructions Executed"]]))
    out = []
    for k in res:
        tot = sum(v for v, _, _ in k) or 1.0
        out.append([(100 * v / tot, s, e) for v, s, e in sorted(k, key=lambda x: -x[0])[:top]])
    return out


def sass_evidence():
    p = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True)
    kern, hist, full = None, collections.defaultdict(collections.Counter), collections.defaultdict(list)
    for line in p.stdout.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            kern = m.group(1)
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if kern and m:
            hist[kern][m.group(2)] += 1
            full[kern].append(line.rstrip())
    return hist, full


def main():
    os.makedirs(OUT, exist_ok=True)
    reps = sys.argv[1:] or sorted(
        os.path.join(ROOT, "gpurun_out", f) for f in os.listdir(os.path.join(ROOT, "gpurun_out")) if f.endswith(".ncu-rep"))
    for rep in reps:
        tag = os.path.basename(rep)[:-8]
        summ = raw_summary(rep)
        st = stalls(rep)
        with open(os.path.join(OUT, tag + ".md"), "w") as f:
            f.write("# ncu --set full summary: %s\n\n(captured with `--clock-control none --import-source on` on a B200; "
                    "durations under ncu are cold-cache and serialised — compare shares, not absolutes)\n\n" % tag)
            for i, d in enumerate(summ):
                f.write("## launch %d: `%s`\n\n| metric | value |\n|---|---|\n" % (i, d["kernel"][:120]))
                for k in WANT:
                    if k in d:
                        f.write("| %s | %s |\n" % (k, d[k]))
                if i < len(st):
                    f.write("\nTop warp-stall samples (share of all samples, SASS, executions):\n\n```\n")
                    for share, src, ex in st[i]:
                        f.write("%5.1f%%  x%-8s %s\n" % (share, ex, src[:110]))
                    f.write("```\n\n")
    hist, full = sass_evidence()
    keys = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "REDG", "ATOMG", "ATOMS", "RED",
            "HMMA", "LDGSTS", "LDG", "STG", "MUFU"]
    with open(os.path.join(OUT, "sass_mnemonics.md"), "w") as f:
        f.write("# SASS evidence per kernel (cuobjdump -sass swiftsnails_b200/lib/libssb200_kernels.so)\n\n"
                "`tcgen05.mma` -> UTC*MMA, `tcgen05.ld` -> LDTM, TMA -> UTMALDG / UBLKCP, vector `red`/`atom` -> REDG / ATOMG\n\n"
                "| kernel | " + " | ".join(keys) + " | total |\n|---|" + "---|" * (len(keys) + 1) + "\n")
        for k in sorted(hist):
            h = hist[k]
            cnt = [sum(v for m, v in h.items() if m.split(".")[0] == key or m.startswith(key + ".")) for key in keys]
            f.write("| `%s` | %s | %d |\n" % (k[:70], " | ".join(str(c) if c else "" for c in cnt), sum(h.values())))
    for k, lines in full.items():
        if "sgns_block_tc" in k:
            with open(os.path.join(OUT, "sgns_block_tc_kernel.sass"), "w") as f:
                f.write("\n".join(lines) + "\n")
    print("profiles written to", OUT)


if __name__ == "__main__":
    main()
