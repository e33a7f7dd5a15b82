"""spmv::cg with the halo push inside the fused step (B200_CG_FUSED_PUSH=1, default) against the separate push
launch (=0): us per iteration on a 27-point 256 x 256 x NZ operator split over the ranks (torchrun), same bits.
  torchrun ... cg_push_probe.py [NZ=256] [stencil=27] [symmetric=0]"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from spmv_b200 import api  # noqa: E402

nz = int(sys.argv[1]) if len(sys.argv) > 1 else 256
stencil = int(sys.argv[2]) if len(sys.argv) > 2 else 27
sym = bool(int(sys.argv[3])) if len(sys.argv) > 3 else False
rank, size, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device(f"cuda:{lr}")
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
api.init_world_from_torch(dist.new_group(backend="gloo"))
ex = api.B200Executor.create(lr)
ex.use_torch_stream()
A = api.create_laplace3d(ex, 256, 256, nz, stencil, sym, api.P2P_NONBLOCKING)
M, cm = A.rows(), A.col_map()
ones = torch.zeros(cm.local_size() + cm.num_ghosts(), dtype=torch.float64, device=dev)
ones[:M] = 1.0
b = torch.empty(M, dtype=torch.float64, device=dev)
cm.update(ones)
A.mult(ones, b)
xs = {m: torch.zeros(M, dtype=torch.float64, device=dev) for m in "01"}
iters = 200
res = {}
for rep in range(3):
    for mode in "01":
        os.environ["B200_CG_FUSED_PUSH"] = mode
        if rep == 0:
            api.cg(ex, A, b, xs[mode], 20, 0.0)  # warm-up: workspace, graph
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        it = api.cg(ex, A, b, xs[mode], iters, 0.0)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        res.setdefault(mode, []).append(ms.item() / it * 1e3)
same = torch.equal(xs["0"], xs["1"])
flag = torch.tensor([1 if same else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
# and a solve to convergence in the default mode
os.environ["B200_CG_FUSED_PUSH"] = "1"
it1, h1 = api.cg(ex, A, b, xs["1"], 5000, 1e-8, history=True)
os.environ["B200_CG_FUSED_PUSH"] = "0"
it0, h0 = api.cg(ex, A, b, xs["0"], 5000, 1e-8, history=True)
if rank == 0:
    print(f"{size} ranks, {stencil}-pt 256x256x{nz} sym={int(sym)}, {M} rows/rank, halo {cm.halo_mode()}: "
          f"separate push {min(res['0']):.1f} us/it  fused push {min(res['1']):.1f} us/it  (runs {res}); "
          f"x bit-identical on every rank: {bool(flag.item())}; solve {it0} / {it1} iterations, "
          f"histories equal: {list(h0) == list(h1)}", flush=True)
A.close()
