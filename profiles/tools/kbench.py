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


def timeit(run, a):
    for _ in range(a.warmup):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / a.iters


def powerlaw(a, dev):
    """--n = rows (config 5 shape, scaled down); general CSR, one block."""
    import numpy as np
    N = a.n
    rp, col, val = synth.powerlaw_rows(N, 0, N)
    rp_d = torch.from_numpy(rp.astype(np.int32)).to(dev)
    ci_d = torch.from_numpy(col.astype(np.int32)).to(dev)
    v_d = torch.from_numpy(val).to(dev)
    nnz = int(rp[-1])
    plan = K.CsrPlan(N, N, rp_d, False, nnz=nnz)
    x = torch.from_numpy(synth.bump_vector(N, 0, N)).to(dev)
    y = torch.empty(N, dtype=torch.float64, device=dev)
    nbytes = 12 * nnz + 4 * (N + 1) + 16 * N
    if a.tune:
        plan.tune(*[int(t) for t in a.tune.split(",")])
    ms = timeit(lambda: K.spmv(plan, rp_d, ci_d, v_d, x, y), a)
    info = plan.info()
    print(f"powerlaw rows={N} nnz={nnz} max_row={info['max_row_len']} tiles={info['num_tiles']} (plan defaults): "
          f"{ms:.4f} ms  {nbytes / ms / 1e6:.1f} GB/s algorithmic", flush=True)
    want = torch.sparse_csr_tensor(rp_d.long(), ci_d.long(), v_d, size=(N, N)).matmul(x)
    print("  rel err vs torch.sparse:", float((y - want).norm() / want.norm()), flush=True)
    if a.tune:
        return
    variants = [int(v) for v in os.environ.get("KB_VARIANTS", "2,3,4,6").split(",")]
    for variant in variants:
        shapes = [(24576, 2, 8, 3), (24576, 2, 16, 2), (49152, 2, 8, 2), (49152, 2, 16, 2), (32768, 2, 8, 3)]
        if variant == 3:
            shapes += [(16384, 2, 16, 2), (16384, 2, 16, 3), (24576, 2, 16, 3), (16384, 3, 16, 3), (24576, 3, 16, 2),
                       (24576, 3, 8, 4), (12288, 2, 16, 3), (12288, 3, 16, 3), (49152, 3, 16, 1), (20480, 2, 16, 3)]
        if variant == 8:  # split walk: A warps products, B warps sums
            shapes = [(24576, 2, 16, 2), (24576, 3, 16, 2), (16384, 3, 16, 2), (16384, 4, 16, 2), (12288, 4, 16, 2),
                      (20480, 3, 16, 2), (32768, 2, 16, 2), (49152, 3, 16, 1), (32768, 4, 16, 1),
                      (16384, 3, 8, 3), (16384, 3, 8, 4), (12288, 4, 8, 4), (8192, 4, 8, 4), (24576, 2, 8, 3),
                      (24576, 3, 8, 3)]
        if variant in (6, 7):  # warp-private pipelined walk: register-bound at 2 CTAs (16 warps) / 3 CTAs (8 warps)
            shapes = [(24576, 2, 16, 2), (24576, 3, 16, 2), (16384, 2, 16, 2), (16384, 3, 16, 2), (16384, 4, 16, 2),
                      (12288, 3, 16, 2), (12288, 4, 16, 2), (8192, 4, 16, 2), (8192, 6, 16, 2), (32768, 3, 16, 2),
                      (49152, 2, 16, 1), (49152, 3, 16, 1), (32768, 4, 16, 1),
                      (16384, 3, 8, 3), (12288, 3, 8, 3), (8192, 4, 8, 3), (24576, 2, 8, 3), (24576, 3, 8, 2)]
        if os.environ.get("KB_SHAPES"):
            shapes = [tuple(int(t) for t in sh.split("x")) for sh in os.environ["KB_SHAPES"].split(",")]
        for tb, stg, cw, ct in shapes:
            plan.tune(variant, tb, stg, cw, ct)
            ms = timeit(lambda: K.spmv(plan, rp_d, ci_d, v_d, x, y), a)
            y.fill_(float("nan"))
            K.spmv(plan, rp_d, ci_d, v_d, x, y)
            err = float((y - want).norm() / want.norm())
            print(f"  variant={variant} tile={tb} stages={stg} cw={cw} ctas={ct}: {ms:.4f} ms  {nbytes / ms / 1e6:.1f} GB/s"
                  f"  err {err:.1e}", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="sym7", choices=["sym7", "sym27", "general7", "general27", "powerlaw"])
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--tune", default="", help="powerlaw: variant,tile_bytes,stages,cwarps,ctas -- time only this shape")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    n = a.n
    if a.case == "powerlaw":
        return powerlaw(a, dev)
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
