import os, sys, torch
sys.path.insert(0, "/root/repo")
os.environ["B200_HOST_TRACE"] = "1"
os.environ["B200_HOST_CHUNKS"] = sys.argv[1] if len(sys.argv) > 1 else "8"
from spmv_b200 import api
torch.cuda.set_device(0)
N = 256**3
xh = torch.randn(N, dtype=torch.float64).pin_memory(); yh = torch.empty(N, dtype=torch.float64).pin_memory()
ex = api.B200Executor.create(0)
mode = os.environ.get("STREAM_MODE", "torch")
if mode == "torch":
    ex.use_torch_stream()
elif mode == "side":
    side = torch.cuda.Stream()
    ex.set_cuda_stream(side.cuda_stream)
print("stream mode", mode, file=sys.stderr)
for st, sym in ((7, True), (27, False)):
    A = api.create_laplace3d(ex, 256, 256, 256, st, sym, api.P2P_BLOCKING)
    for _ in range(3):
        A.mult_host(xh.numpy(), yh.numpy())
    A.close()
