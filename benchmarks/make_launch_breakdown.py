#!/usr/bin/env python
"""profiles/launch_breakdown.md from an `ncu --metrics gpu__time_duration.sum --csv` log of
`bench.py --pipeline 0`: every kernel launched by the word2vec steps with its device time, aggregated
per kernel.  Launches before the first step (table set-up, pre-population) are left out: the first
step starts one launch before the first `neg_sample_kernel` (`build_keys_kernel`).
usage: python benchmarks/make_launch_breakdown.py gpurun_out/launches.csv [tokens_per_step]"""
import collections
import csv
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    path = sys.argv[1]
    tps = sys.argv[2] if len(sys.argv) > 2 else "524 288"
    lines = [l for l in open(path) if l.startswith('"')]
    rows = list(csv.DictReader(lines))
    launches = [(r["Kernel Name"], float(r["Metric Value"]) / 1e3) for r in rows if r["Metric Name"] == "gpu__time_duration.sum"]
    first = next(i for i, (k, _) in enumerate(launches) if "neg_sample_kernel" in k)
    step = launches[max(0, first - 1):]          # the step starts with build_keys_kernel, one launch before the sampler
    n_steps = sum(1 for k, _ in step if "sgns_block_tc_kernel" in k or "sgns_tiles_kernel" in k)
    agg = collections.OrderedDict()
    for k, us in step:
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += us
    total = sum(a[1] for a in agg.values())
    out = os.path.join(ROOT, "profiles", "launch_breakdown.md")
    with open(out, "w") as f:
        f.write("# per-kernel device time of the word2vec step (ncu --metrics gpu__time_duration.sum, 1 B200, non-pipelined)\n")
        f.write("%s tokens/step, vocab 1M, D=128, AdaGrad, pre-populated table, %d steps captured (%d launches, %.1f us of kernels per step); "
                "ncu serialises launches and flushes caches: compare shares.\n\n" % (tps, n_steps, len(step), total / max(n_steps, 1)))
        f.write("| kernel | launches | us / launch | us / step | share |\n|---|---|---|---|---|\n")
        for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("| `%s` | %d | %.1f | %.1f | %.1f%% |\n" % (k[:60], n, us / n, us / max(n_steps, 1), 100 * us / total))
    print("written", out, "steps", n_steps, "launches", len(step))


if __name__ == "__main__":
    main()
