#!/usr/bin/env python
"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list of one bench step into per-kernel shares
(the per-launch times are cold-cache and serialised under the profiler: the SHARE of each kernel is what is compared
with bench.py's live CUDA-event numbers).   python tools/launch_shares.py gpurun_out/X_launches.csv profiles/X_launch_shares.txt"""
import collections
import csv
import re
import sys


def main():
    src, dst = sys.argv[1], sys.argv[2]
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"')) if len(r) > 14]
    hdr, rows = rows[0], rows[1:]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot = collections.OrderedDict()
    for r in rows:
        name = re.sub(r"\(.*", "", r[kn].replace("void ", "").replace("ts2d::", ""))
        name = re.sub(r"<.*", "", name) if not name.startswith(("conv_", "project", "blend")) else re.sub(r"__half", "fp16", re.sub(r"__nv_bfloat16", "bf16", name))
        ns = float(r[mv].replace(",", "")) * {"ns": 1.0, "us": 1e3, "ms": 1e6}.get(r[mu], 1.0)
        n, t = tot.get(name, (0, 0.0))
        tot[name] = (n + 1, t + ns)
    total = sum(t for _, t in tot.values())
    lines = ["# %s: %d launches, %.3f ms summed kernel time under ncu (cold-cache, serialised)" % (src.split("/")[-1], len(rows), total / 1e6),
             "%-60s %7s %10s %7s" % ("kernel", "count", "ms", "share")]
    for name, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        lines.append("%-60s %7d %10.3f %6.1f%%" % (name[:60], n, t / 1e6, 100.0 * t / total))
    open(dst, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
