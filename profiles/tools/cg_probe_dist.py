"""CG solve time with the peer-memory scalar allreduce vs NCCL (torchrun; strong-scaled 27-point 256^3)."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from spmv_b200 import api  # noqa: E402

rank, size, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device(f"cuda:{lr}")
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
api.init_world_from_torch(dist.new_group(backend="gloo"))
ex = api.B200Executor.create(lr)
ex.use_torch_stream()
mats = {}
for mode in ("peer", "nccl"):
    os.environ["B200_ALLREDUCE"] = mode
    mats[mode] = api.create_laplace3d(ex, 256, 256, 256, 27, False, api.P2P_NONBLOCKING)
A = mats["peer"]
M, cm = A.rows(), A.col_map()
ones = torch.zeros(cm.local_size() + cm.num_ghosts(), dtype=torch.float64, device=dev)
ones[:M] = 1.0
b = torch.empty(M, dtype=torch.float64, device=dev)
cm.update(ones)
A.mult(ones, b)
xs = torch.zeros(M, dtype=torch.float64, device=dev)
for rep in range(4):
    for mode in ("peer", "nccl"):
        dist.barrier()
        torch.cuda.synchronize()
        t = time.perf_counter()
        it = api.cg(ex, mats[mode], b, xs, 5000, 1e-8)
        torch.cuda.synchronize()
        ms = torch.tensor([(time.perf_counter() - t) * 1e3], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"{size} ranks {mode}: {it} iterations {ms.item():8.2f} ms  {ms.item() / it * 1e3:7.1f} us/it", flush=True)
for m in mats.values():
    m.close()
