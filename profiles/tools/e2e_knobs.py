import os, sys, time, torch
sys.path.insert(0, "/root/repo")
from spmv_b200 import api
torch.cuda.set_device(0)
N = 256**3
xh = torch.randn(N, dtype=torch.float64).pin_memory(); yh = torch.empty(N, dtype=torch.float64).pin_memory()
ex = api.B200Executor.create(0); ex.use_torch_stream()
for st, sym in ((7, True), (27, False)):
    A = api.create_laplace3d(ex, 256, 256, 256, st, sym, api.P2P_BLOCKING)
    for _ in range(3):
        A.mult_host(xh.numpy(), yh.numpy())
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(10):
        A.mult_host(xh.numpy(), yh.numpy())
    print(st, sym, os.environ.get("B200_HOST_CHUNKS"), os.environ.get("B200_TILES_GRID"), os.environ.get("B200_HOST_SKIP_KERNEL"), f"{(time.perf_counter()-t)*100:.3f} ms")
    A.close()
