#!/usr/bin/env python
"""za_probe.py -- sweeper for the symmetric SpMV: zero-ahead of y (B200_SPMV_ZA) against the memset pass and
launch shapes.  ZA_CONFIGS = comma list of za:k:d[:stages[:tile_bytes[:cwarps[:ctas]]]] (k tiles per chunk, d chunks
of lead); every config is timed, compared with the first one and run on NaN-poisoned y (coverage of the zero-fill).
ZA_GCONFIGS = stages:tile:cw:ctas shapes for the general-storage kernel of the same stencil.
The za / k / d fields only act on a build with profiles/r01_exp_zero_ahead_lead.patch applied (the product has no
zero-ahead path: measured slower, profiles/r01_sym7_supply.md); on the product build they are ignored and the tool
is a launch-shape sweeper."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from spmv_b200 import kernels as K  # noqa: E402
from spmv_b200 import synth  # noqa: E402


def timeit(run, iters=50, warmup=5):
    for _ in range(warmup):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    dev = torch.device("cuda:0")
    n = int(os.environ.get("ZA_N", "256"))
    N = n ** 3
    x = torch.from_numpy(synth.bump_vector(N, 0, N)).to(dev)
    for pts in [int(t) for t in os.environ.get("ZA_PTS", "7,27").split(",")]:
        s = K.laplace3d(n, n, n, 0, n, pts, 1, dev)
        rp, ci, v = s["local"]
        nnz = int(ci.numel())
        nbytes = 12 * nnz + 4 * (N + 1) + 24 * N
        res = {}
        configs = [c.split(":") for c in os.environ.get("ZA_CONFIGS", "0:4:2,1:4:2,1:4:4").split(",")]
        for cfg in configs:
            za, k, d = cfg[:3]
            os.environ["B200_SPMV_ZA"] = za
            os.environ["B200_SPMV_ZA_K"] = k
            os.environ["B200_SPMV_ZA_D"] = d
            for name, val in zip(("STAGES", "TILE_BYTES", "CWARPS", "CTAS"), list(cfg[3:7]) + [""] * 4):
                if val:
                    os.environ["B200_SPMV_" + name] = val
                else:
                    os.environ.pop("B200_SPMV_" + name, None)
            plan = K.CsrPlan(N, N, rp, True)
            y = torch.empty(N, dtype=torch.float64, device=dev)
            bad = 0
            for rep in range(10):
                y.fill_(float("nan"))
                K.symv(plan, rp, ci, v, s["diag"], x, y)
                torch.cuda.synchronize()
                if "ref" not in res:
                    res["ref"] = y.clone()
                err = float((y - res["ref"]).norm() / res["ref"].norm())
                if not err < 1e-13:
                    bad += 1
            ms = timeit(lambda: K.symv(plan, rp, ci, v, s["diag"], x, y))
            print(f"sym{pts} n={n} cfg={':'.join(cfg)} tiles={plan.info()['num_tiles']}: {ms:.4f} ms  "
                  f"{nbytes / ms / 1e6:.1f} GB/s  last err {err:.2e} bad {bad}/10", flush=True)
            del plan
        for name in ("STAGES", "TILE_BYTES", "CWARPS", "CTAS"):
            os.environ.pop("B200_SPMV_" + name, None)
        # general-storage cross-check of the reference result
        g = K.laplace3d(n, n, n, 0, n, pts, 0, dev)
        grp, gci, gv = g["local"]
        gplan = K.CsrPlan(N, N, grp, False)
        yg = torch.empty(N, dtype=torch.float64, device=dev)
        K.spmv(gplan, grp, gci, gv, x, yg)
        print(f"  sym vs general storage: {float((res['ref'] - yg).norm() / yg.norm()):.2e}", flush=True)
        gb = 12 * int(gci.numel()) + 4 * (N + 1) + 16 * N
        ms = timeit(lambda: K.spmv(gplan, grp, gci, gv, x, yg))
        print(f"general{pts} n={n}: {ms:.4f} ms  {gb / ms / 1e6:.1f} GB/s", flush=True)
        for name in ("STAGES", "TILE_BYTES", "CWARPS", "CTAS"):
            os.environ.pop("B200_SPMV_" + name, None)
        for gcfg in [c for c in os.environ.get("ZA_GCONFIGS", "").split(",") if c]:
            st, tb, cw, ct = [int(t) for t in gcfg.split(":")]
            try:
                gplan.tune(0, tb, st, cw, ct)
                ms = timeit(lambda: K.spmv(gplan, grp, gci, gv, x, yg))
                print(f"general{pts} n={n} stages={st} tile={tb} cw={cw} ctas={ct}: {ms:.4f} ms  {gb / ms / 1e6:.1f} GB/s", flush=True)
            except Exception as e:  # shapes that do not fit shared memory
                print(f"general{pts} {gcfg}: {str(e)[:80]}", flush=True)
        del g, gplan, s


if __name__ == "__main__":
    main()
