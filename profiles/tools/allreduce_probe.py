"""Scalar allreduce latency, peer-memory kernel vs NCCL (torchrun, one rank per GPU)."""
import os
import sys

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
A = api.create_laplace3d(ex, 32, 32, 8 * size, 7, False, api.P2P_NONBLOCKING)
cm = A.col_map()
v = torch.zeros(1, dtype=torch.float64, device=dev)
work = torch.zeros(1 << 22, dtype=torch.float64, device=dev)
for nccl in (False, True, False, True):
    for with_work in (False, True):
        def step():
            if with_work:
                work.add_(1.0)  # a 32 MB kernel between the reductions, as in the CG loop
            cm.allreduce_sum(v, nccl=nccl)
        v.fill_(rank + 1.0)
        cm.allreduce_sum(v, nccl=nccl)
        torch.cuda.synchronize()
        assert v.item() == size * (size + 1) / 2, v.item()
        for _ in range(20):
            step()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 500
        e0.record()
        for _ in range(n):
            step()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / n * 1e3
        if with_work:
            e0.record()
            for _ in range(n):
                work.add_(1.0)
            e1.record()
            torch.cuda.synchronize()
            us -= e0.elapsed_time(e1) / n * 1e3
        t = torch.tensor([us], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"{size} ranks  {'nccl' if nccl else 'peer'}  {'between kernels' if with_work else 'back to back  '}: {t.item():7.2f} us per allreduce", flush=True)
A.close()
