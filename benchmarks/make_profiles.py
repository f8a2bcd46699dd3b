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


def base_name(kernel):
    """'void ssb::push_kernel<1>(SsbTableView, ...)' -> 'push_kernel'"""
    k = kernel.split("(")[0]
    k = re.sub(r"<.*", "", k).strip()
    return k.split("::")[-1].split(" ")[-1]


def stalls_by_source_line(rep, kernel, top=22):
    """Warp-stall samples of one kernel aggregated per CUDA source line (needs -lineinfo + --import-source on):
    [(share %, 'file:line', source text, 'reason=n, reason=n')]"""
    p = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                        "regex:" + kernel], capture_output=True, text=True)
    rows = list(csv.reader(io.StringIO(p.stdout)))
    fname, hdr, agg = None, None, {}
    for r in rows:
        if r and r[0] == "File Path":
            fname = os.path.basename(r[1])
            continue
        if r and r[0] == "Line No":
            hdr = r
            continue
        if not hdr or len(r) < 8 or r[2] != "-" or not r[0].isdigit():
            continue
        try:
            n = int(r[hdr.index("# Samples")])
        except ValueError:
            continue
        if n == 0:
            continue
        key = (fname, int(r[0]))
        ent = agg.setdefault(key, [0, r[1], collections.Counter()])
        ent[0] += n
        for i, h in enumerate(hdr):
            if h.startswith("stall_") and "Not Issued" not in h and i < len(r) and r[i].isdigit():
                ent[2][h[6:]] += int(r[i])
    tot = sum(e[0] for e in agg.values()) or 1
    out = []
    for (f, ln), (n, src, reasons) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        out.append((100.0 * n / tot, "%s:%d" % (f, ln), " ".join(src.split())[:96],
                    ", ".join("%s=%d" % (k, v) for k, v in reasons.most_common(2))))
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
        st_cache = {}
        with open(os.path.join(OUT, tag + ".md"), "w") as f:
            f.write("# ncu --set full summary: %s\n\n(captured with `--clock-control none --import-source on` on a B200; "
                    "durations under ncu are cold-cache and serialised — compare shares, not absolutes)\n\n" % tag)
            for i, d in enumerate(summ):
                f.write("## launch %d: `%s`\n\n| metric | value |\n|---|---|\n" % (i, d["kernel"][:120]))
                for k in WANT:
                    if k in d:
                        f.write("| %s | %s |\n" % (k, d[k]))
                bn = base_name(d["kernel"])
                if bn not in st_cache:
                    st_cache[bn] = stalls_by_source_line(rep, bn)
                    if st_cache[bn]:
                        f.write("\nWarp-stall samples per CUDA source line (share of the kernel's samples; top two stall reasons; "
                                "all launches of this kernel in the report):\n\n```\n")
                        for share, loc, src, why in st_cache[bn]:
                            f.write("%5.1f%%  %-18s %-96s [%s]\n" % (share, loc, src, why))
                        f.write("```\n\n")
                else:
                    f.write("\n(stall samples: see the first launch of this kernel above)\n\n")
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
    # one SASS listing per kernel (cuobjdump text, encodings already stripped by the regex above)
    sass_dir = os.path.join(OUT, "sass")
    os.makedirs(sass_dir, exist_ok=True)
    for old in os.listdir(sass_dir):
        os.remove(os.path.join(sass_dir, old))
    dem = subprocess.run(["cu++filt"] + list(full), capture_output=True, text=True).stdout.splitlines()
    for (k, lines), d in zip(full.items(), dem if len(dem) == len(full) else list(full)):
        name = re.sub(r"\(.*", "", re.sub(r"\((bool|int|unsigned int)\)", "", d)).replace("void ", "").replace("ssb::", "").replace("::", "_")
        name = re.sub(r"[<>, ]+", "_", name).strip("_")
        with open(os.path.join(sass_dir, name + ".sass"), "w") as f:
            f.write("\t\tFunction : %s\n" % k + "\n".join(lines) + "\n")
        if "tc20sgns_block_tc_kernelILb0E" in k:      # the production instantiation of the pipelined kernel
            with open(os.path.join(OUT, "sgns_block_tc_kernel.sass"), "w") as f:
                f.write("\n".join(lines) + "\n")
    print("profiles written to", OUT)


if __name__ == "__main__":
    main()
