"""PCIe ceiling for the host-vector SpMV and the chunk-count sweep of Matrix::mult_host.
usage: python profiles/tools/e2e_probe.py   (one GPU)"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from spmv_b200 import api  # noqa: E402

N = 256 ** 3
dev = torch.device("cuda:0")
torch.cuda.set_device(0)
xh = torch.randn(N, dtype=torch.float64).pin_memory()
yh = torch.empty(N, dtype=torch.float64).pin_memory()
xd = torch.empty(N, dtype=torch.float64, device=dev)
yd = torch.randn(N, dtype=torch.float64, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def wall(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / n * 1e3


def h2d():
    xd.copy_(xh, non_blocking=True)


def d2h():
    yh.copy_(yd, non_blocking=True)


def both():
    with torch.cuda.stream(s1):
        xd.copy_(xh, non_blocking=True)
    with torch.cuda.stream(s2):
        yh.copy_(yd, non_blocking=True)


mb = N * 8 / 1e6
for name, fn in (("h2d alone", h2d), ("d2h alone", d2h), ("h2d + d2h concurrently", both)):
    ms = wall(fn)
    print(f"{name:28s} {ms:7.3f} ms  {mb / ms:6.1f} GB/s per direction")

ex = api.B200Executor.create(0)
ex.use_torch_stream()
for stencil, sym in ((7, True), (27, False)):
    for chunks in (1, 2, 4, 8, 16, 32, 64, 128):
        os.environ["B200_HOST_CHUNKS"] = str(chunks)
        A = api.create_laplace3d(ex, 256, 256, 256, stencil, sym, api.P2P_BLOCKING)
        ms = wall(lambda: A.mult_host(xh.numpy(), yh.numpy()))
        print(f"mult_host {stencil}-pt sym={sym} chunks={chunks:4d}: {ms:7.3f} ms")
        A.close()
