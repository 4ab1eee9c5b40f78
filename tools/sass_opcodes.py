#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of the shipped library: the evidence that the hot kernels are Blackwell-native
(`tcgen05.mma` -> UTC*MMA, `tcgen05.ld` -> LDTM, TMA -> UTMALDG / UBLKCP, mbarrier -> SYNCS, `tcgen05.commit` -> UTCBAR)
and not recompiled `mma.sync` code (HMMA).  Runs on the CPU box: needs cuobjdump and c++filt only.

    python tools/sass_opcodes.py [path/to/libts2d_b200.so] > profiles/r02_sass_opcodes.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WATCH = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCATOM", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "SYNCS", "HMMA", "LDGSTS",
         "FADD2", "FFMA2", "FMUL2", "STG.E.ENL2.256", "LDG.E.ENL2.256", "LDG.E.128", "STG.E.128", "ATOMG", "RED", "ACQBULK", "UCGABAR", "ELECT", "IDP"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "totalsegmentator2d_b200", "libts2d_b200.so")
    sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    cur = None
    op_re = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)")
    for line in sass.splitlines():
        if "Function :" in line:
            cur = line.split("Function :")[1].strip()
            per[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = op_re.match(line)
        if m:
            per[cur][m.group(1)] += 1
    dm = demangle(list(per))
    print(f"# SASS opcode histogram of {os.path.relpath(so, ROOT)} (cuobjdump -sass, sm_100a); {len(per)} kernels")
    print("# columns: instructions | tensor-core / TMEM / TMA / barrier opcodes (count) ... ; kernels without any of them are listed at the end")
    tot = collections.Counter()
    rest = []
    for k, cnt in per.items():
        n = sum(cnt.values())
        hits = []
        for w in WATCH:
            c = sum(v for op, v in cnt.items() if op == w or op.startswith(w + ".") or (w.count(".") and op.startswith(w)))
            if c:
                hits.append(f"{w} {c}")
                tot[w] += c
        name = re.sub(r"\(.*$", "", dm[k]).replace("void ", "")
        if any(h.split()[0] in ("UTCHMMA", "UTCQMMA", "LDTM", "UTMALDG", "UBLKCP", "UTCBAR") for h in hits):
            print(f"{name:<95s} {n:6d} | " + ", ".join(hits))
        else:
            rest.append((name, n, hits))
    print("# ---- kernels without tcgen05 / TMA opcodes (CUDA-core kernels: HBM-bound stages, parity locks) ----")
    for name, n, hits in rest:
        print(f"{name:<95s} {n:6d} | " + ", ".join(hits))
    print("# ---- totals ----")
    print("# " + ", ".join(f"{w} {c}" for w, c in tot.items()))
    if tot.get("HMMA"):
        print("# WARNING: legacy HMMA (mma.sync) instructions present")


if __name__ == "__main__":
    main()
