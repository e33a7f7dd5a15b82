#!/usr/bin/env python
"""kbench.py -- single-kernel timing / ncu target for the SpMV kernels (tuning tool, not the bench).

  python profiles/tools/kbench.py --case sym7 --n 256 --iters 20
  B200_SPMV_* environment variables select the plan tuning (see b200_csr_plan_create).
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from spmv_b200 import kernels as K  # noqa: E402
from spmv_b200 import synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="sym7", choices=["sym7", "sym27", "general7", "general27"])
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    n = a.n
    N = n * n * n
    sym = a.case.startswith("sym")
    stencil = 7 if a.case.endswith("7") and not a.case.endswith("27") else 27
    g = K.laplace3d(n, n, n, 0, n, stencil, 1 if sym else 0, dev)
    rp, ci, v = g["local"]
    nnz = int(rp[-1].item())
    plan = K.CsrPlan(N, N, rp, sym, nnz=nnz)
    x = torch.from_numpy(synth.bump_vector(N, 0, N)).to(dev)
    y = torch.empty(N, dtype=torch.float64, device=dev)
    if sym:
        run = lambda: K.symv(plan, rp, ci, v, g["diag"], x, y)
        nbytes = 12 * nnz + 4 * (N + 1) + 24 * N
    else:
        run = lambda: K.spmv(plan, rp, ci, v, x, y)
        nbytes = 12 * nnz + 4 * (N + 1) + 16 * N
    for _ in range(a.warmup):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    print(f"{a.case} n={n} nnz={nnz} tiles={plan.info()['num_tiles']}: {ms:.4f} ms  {nbytes / ms / 1e6:.1f} GB/s algorithmic "
          f"env={ {k: v for k, v in os.environ.items() if k.startswith('B200_')} }", flush=True)


if __name__ == "__main__":
    main()
