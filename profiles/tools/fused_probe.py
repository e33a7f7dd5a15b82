#!/usr/bin/env python
"""fused_probe.py -- vector step of spmv::cg on a strong-scaling-sized slab (27-point 256x256xNZ on one GPU), separate
kernels against the fused cooperative step at several grid sizes.  B200_CG_FUSED_CTAS is read once per process, so the
sweep re-executes itself.   python profiles/tools/fused_probe.py [nz]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def one(nz):
    import torch
    from spmv_b200 import api
    torch.cuda.set_device(0)
    ex = api.B200Executor.create(0)
    ex.use_torch_stream()
    A = api.create_laplace3d(ex, 256, 256, nz, 27, False, api.P2P_BLOCKING)
    M = A.rows()
    ones = torch.ones(M, dtype=torch.float64, device="cuda")
    b = torch.empty_like(ones)
    A.mult(ones, b)
    x = torch.zeros_like(b)
    api.cg(ex, A, b, x, 10, 0.0)
    _, ph = api.cg_phases(ex, A, b, x, 40, 0.0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    it = api.cg(ex, A, b, x, 200, 0.0)
    e1.record()
    torch.cuda.synchronize()
    print(f"nz={nz} rows={M} FUSED_VEC={os.environ.get('B200_CG_FUSED_VEC', '-')} CTAS={os.environ.get('B200_CG_FUSED_CTAS', '-')} "
          f"FUSED_DOT={os.environ.get('B200_CG_FUSED_DOT', '-')}: local {ph['local']:.1f} us, vector {ph['vector']:.1f} us, "
          f"graph path {e0.elapsed_time(e1) / it * 1e3:.1f} us/iter", flush=True)


if __name__ == "__main__":
    if os.environ.get("FUSED_PROBE_CHILD"):
        one(int(sys.argv[1]))
    else:
        nz = sys.argv[1] if len(sys.argv) > 1 else "32"
        for env in ({"B200_CG_FUSED_VEC": "0"}, {"B200_CG_FUSED_VEC": "0", "B200_CG_FUSED_DOT": "0"},
                    {"B200_CG_FUSED_VEC": "1", "B200_CG_FUSED_CTAS": "1"}, {"B200_CG_FUSED_VEC": "1", "B200_CG_FUSED_CTAS": "2"},
                    {"B200_CG_FUSED_VEC": "1", "B200_CG_FUSED_CTAS": "3"}, {"B200_CG_FUSED_VEC": "1", "B200_CG_FUSED_CTAS": "6"},
                    {"B200_CG_FUSED_VEC": "1", "B200_CG_FUSED_CTAS": "2", "B200_CG_FUSED_DOT": "0"},
                    {"B200_CG_FUSED_VEC": "1", "B200_CG_FUSED_CTAS": "6", "B200_CG_FUSED_DOT": "0"}):
            subprocess.run([sys.executable, __file__, nz], env=dict(os.environ, FUSED_PROBE_CHILD="1", **env))
