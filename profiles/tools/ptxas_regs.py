#!/usr/bin/env python
"""Registers / spills per kernel from the build's `-Xptxas -v` logs (build/kernels/*.ptxas.log).
  python profiles/tools/ptxas_regs.py [substring ...]     # keep kernels whose demangled name has every substring
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for f in sorted(os.listdir(os.path.join(ROOT, "build", "kernels"))):
    if not f.endswith(".ptxas.log"):
        continue
    log = open(os.path.join(ROOT, "build", "kernels", f)).read()
    blocks = re.split(r"ptxas info\s+: Compiling entry function '([^']+)' for 'sm_100a'", log)
    for i in range(1, len(blocks), 2):
        dem = subprocess.run(["c++filt", blocks[i]], capture_output=True, text=True).stdout.strip()
        if not all(k in dem for k in sys.argv[1:]):
            continue
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", blocks[i + 1])
        r = re.search(r"Used (\d+) registers", blocks[i + 1])
        print(f"{dem[:110]:110s} regs {r.group(1):>3s} stack {m.group(1):>4s} spill st/ld {m.group(2)}/{m.group(3)}")
