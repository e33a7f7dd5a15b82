#!/usr/bin/env python
"""sass_histogram.py -- per-kernel instruction histogram of spmv_b200/lib/libb200spmv.so (cuobjdump -sass), so
the Blackwell-native claims (bulk-async copies UBLKCP, mbarrier SYNCS, fire-and-forget REDG.E.ADD.F64, no
library kernels) can be checked without the (git-ignored) binary.
  python profiles/tools/sass_histogram.py > profiles/r02_sass_histogram.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB = os.path.join(ROOT, "spmv_b200", "lib", "libb200spmv.so")
KEYS = ["UBLKCP", "SYNCS", "REDG", "ATOMG", "LDG", "STG", "LDS", "STS", "DFMA", "DMUL", "DADD", "BAR", "SHFL",
        "FENCE", "MEMBAR", "CCTL", "ERRBAR", "LD.", "ST."]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", out)))
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            op = m.group(1)
            kernels[cur]["_total"] += 1
            kernels[cur][op.split(".")[0]] += 1
            if op.startswith("REDG") or op.startswith("UBLKCP") or op.startswith("SYNCS") or op.startswith("ATOMG"):
                kernels[cur][op] += 1
    names = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    print(f"# {os.path.relpath(LIB, ROOT)}: {len(kernels)} kernels, cubin architectures {arch}")
    tot = collections.Counter()
    for (mangled, c), name in zip(kernels.items(), names):
        for k, v in c.items():
            tot[k] += v
    print("# whole library: " + ", ".join(f"{k} {tot[k]}" for k in ("_total", "UBLKCP", "SYNCS", "REDG", "ATOMG", "LDG", "LDS", "STS", "BAR", "SHFL") if tot[k]))
    full = sorted(k for k in tot if "." in k)
    print("# full mnemonics: " + ", ".join(f"{k} {tot[k]}" for k in full))
    print()
    want = sys.argv[1:] or ["spmv_pipe_kernel<double", "spmv_irr_kernel<double", "halo_", "cg_fused", "update_", "dot_kernel", "peer_", "gather", "scatter", "zero_fill", "rowlist"]
    for (mangled, c), name in zip(kernels.items(), names):
        short = re.sub(r"\(.*", "", name).replace("void b200::", "").replace("b200::", "")
        if not any(w in short for w in want):
            continue
        cols = " ".join(f"{k}={c[k]}" for k in ("UBLKCP", "SYNCS", "REDG", "ATOMG", "LDG", "STG", "LDS", "STS", "DADD", "DMUL", "DFMA", "BAR", "SHFL") if c[k])
        print(f"{short[:100]:100s} n={c['_total']:5d}  {cols}")


if __name__ == "__main__":
    main()
