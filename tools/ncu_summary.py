#!/usr/bin/env python
"""Condenses `ncu -i X.ncu-rep --page raw --csv` of one forward's conv launches (tools/gpu_round_check.sh) into the
table committed under profiles/ and into profiles/ncu_traffic.json, which bench.py reads for `roofline.traffic`
(DRAM bytes of the conv launches of one study, measured by ncu; per-launch times under ncu are cold-cache and
serialised, so only byte counts and shares are taken from here, never durations).

    python tools/ncu_summary.py gpurun_out/r01_convs_raw.csv profiles/r01_ncu_convs_summary.txt [--batch 12]
"""
import argparse
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

COLS = {
    "us": "gpu__time_duration.sum",
    "rd": "dram__bytes_read.sum",
    "wr": "dram__bytes_write.sum",
    "grid": "launch__grid_size",
    "tensor_pct": "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
    "tmem_pct": "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "dram_pct": "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts_pct": "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "regs": "launch__registers_per_thread",
    "smem": "launch__shared_mem_per_block_dynamic",
}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}

# launch order of one CS-512 forward (37 conv / transposed-conv launches; norm passes are not captured by -k regex:conv_)
NAMES = (["enc%d.%d" % (s, i) for s in range(8) for i in range(2)] +
         [n for d in range(7) for n in ("up%d" % d, "dec%d.0" % d, "dec%d.1" % d)])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("raw_csv")
    ap.add_argument("out_txt")
    ap.add_argument("--batch", type=int, default=12)
    ap.add_argument("--json", default=os.path.join(ROOT, "profiles", "ncu_traffic.json"))
    a = ap.parse_args()
    rows = list(csv.reader(open(a.raw_csv)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {}
    for k, name in COLS.items():
        idx[k] = hdr.index(name) if name in hdr else -1
    if idx["tensor_pct"] < 0 or all(r[idx["tensor_pct"]] in ("", "n/a", "0") for r in data):
        # fp16 operands are counted by their own tensor-pipe metric
        alt = COLS["tensor_pct"].replace("src_bf16", "src_fp16")
        if alt in hdr:
            idx["tensor_pct"] = hdr.index(alt)
    kn = hdr.index("Kernel Name")

    def val(r, k):
        i = idx[k]
        if i < 0 or r[i] in ("", "n/a"):
            return float("nan")
        return float(r[i].replace(",", "")) * UNIT.get(units[i], 1.0)

    lines = ["# ncu --set full --clock-control none, one CS-512 forward at batch %d (second forward of tools/ncu_forward.py);" % a.batch,
             "# durations are cold-cache / serialised under the profiler: use shares and byte counts, not absolute times",
             "%-8s %-18s %8s %6s %9s %9s %7s %7s %7s %7s %5s" % ("layer", "kernel", "us", "grid", "dramRdMB", "dramWrMB", "tens%", "tmem%", "dram%", "lts%", "regs")]
    tot_us = tot_b = 0.0
    per = []
    for j, r in enumerate(data):
        name = NAMES[j] if len(data) == len(NAMES) else str(j)
        k = r[kn].replace("void ", "").replace("ts2d::", "")
        k = k[:k.index("(")] if "(" in k else k
        us, rd, wr = val(r, "us"), val(r, "rd"), val(r, "wr")
        tot_us += us
        tot_b += rd + wr
        per.append({"layer": name, "kernel": k, "us": us, "dram_read_bytes": rd, "dram_write_bytes": wr})
        lines.append("%-8s %-18s %8.1f %6d %9.1f %9.1f %7.2f %7.2f %7.2f %7.2f %5d" % (
            name, k[:18], us, int(val(r, "grid")), rd / 1e6, wr / 1e6, val(r, "tensor_pct"), val(r, "tmem_pct"), val(r, "dram_pct"), val(r, "lts_pct"),
            int(val(r, "regs"))))
    lines.append("# total: %.1f us under ncu, DRAM read+write %.1f MB per forward of %d tiles" % (tot_us, tot_b / 1e6, a.batch))
    open(a.out_txt, "w").write("\n".join(lines) + "\n")
    keep = {}
    try:
        keep = {k: v for k, v in json.load(open(a.json)).items() if k not in ("source", "batch", "conv_dram_bytes_per_forward", "launches")}
    except Exception:
        keep = {}
    out = {"source": os.path.basename(a.raw_csv), "batch": a.batch, "conv_dram_bytes_per_forward": tot_b, "launches": per}
    out.update(keep)        # other kernels' captures (e.g. project_strided_i16_kernel) live in the same file
    json.dump(out, open(a.json, "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    sys.exit(main())
