"""Fixed cost of one spmv::cg call: time consecutive calls with different kmax (one GPU)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from spmv_b200 import api  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda:0")
ex = api.B200Executor.create(0)
ex.use_torch_stream()
for st, sym in ((27, False), (7, True), (7, False)):
    A = api.create_laplace3d(ex, 256, 256, 256, st, sym, api.P2P_BLOCKING)
    M = A.rows()
    ones = torch.ones(M, dtype=torch.float64, device=dev)
    b = torch.empty(M, dtype=torch.float64, device=dev)
    A.mult(ones, b)
    xs = torch.zeros(M, dtype=torch.float64, device=dev)
    if os.environ.get("WITH_HOST"):
        xh = torch.randn(M, dtype=torch.float64).pin_memory()
        yh = torch.empty(M, dtype=torch.float64).pin_memory()
        for _ in range(3):
            A.mult_host(xh.numpy(), yh.numpy())
        if os.environ.get("WITH_HOST") == "2":
            del xh, yh
    for kmax, rtol in ((10, 0.0), (100, 0.0), (100, 0.0), (200, 0.0), (100, 1e-30), (50, 0.0), (100, 0.0)):
        torch.cuda.synchronize()
        t = time.perf_counter()
        it = api.cg(ex, A, b, xs, kmax, rtol)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t) * 1e3
        print(f"{st}-pt sym={sym} kmax={kmax:4d} rtol={rtol:g}: {it:4d} iterations {ms:8.2f} ms  {ms / max(it, 1):6.3f} ms/it")
    A.close()
