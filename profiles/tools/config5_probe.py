#!/usr/bin/env python
"""config5_probe.py -- where the config-5 step goes at N ranks (torchrun): halo alone, mult alone,
each block's kernel alone (kernel-level API on copies of the blocks), pack alone.  Prints per rank."""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from spmv_b200 import api, synth  # noqa: E402
from spmv_b200 import kernels as K  # noqa: E402


def ev():
    return torch.cuda.Event(enable_timing=True)


def timeit(fn, n=30, sync=None):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    if sync:
        sync()
    e0, e1 = ev(), ev()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def main():
    rank, size, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    rows_per_gpu = int(sys.argv[1]) if len(sys.argv) > 1 else 6_250_000
    dev = torch.device("cuda", lr)
    torch.cuda.set_device(dev)
    bar = lambda: None  # noqa: E731
    if size > 1:
        dist.init_process_group("nccl", device_id=dev)
        api.init_world_from_torch(dist.new_group(backend="gloo"))
        bar = dist.barrier
    ex = api.B200Executor.create(lr)
    ex.use_torch_stream()
    n = rows_per_gpu * size
    path = os.path.join("/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir(), f"c5probe_{os.getpid() if size == 1 else os.environ['MASTER_PORT']}.petsc")
    if rank == 0:
        api.write_powerlaw_petsc(path, n)
    bar()
    A = api.read_petsc_binary_matrix(path, ex, False, api.P2P_NONBLOCKING if size > 1 else api.P2P_BLOCKING)
    bar()
    if rank == 0:
        os.remove(path)
    cm = A.col_map()
    rows, ng = A.rows(), cm.num_ghosts()
    r0 = int(synth.owner_ranges(size, n)[rank])
    x = torch.zeros(rows + ng, dtype=torch.float64, device=dev)
    x[:rows] = torch.from_numpy(synth.bump_vector(n, r0, r0 + rows)).to(dev)
    y = torch.empty(rows, dtype=torch.float64, device=dev)
    out = [f"[rank {rank}/{size}] rows {rows} ghosts {ng} neighbours {len(cm.get('neighbours')) if size > 1 else 0}"]
    t = timeit(lambda: (cm.update(x), A.mult(x, y)), sync=bar)
    out.append(f"update+mult {t:.1f} us")
    if size > 1:
        t = timeit(lambda: (cm.update(x), cm.update_finalise(x)), sync=bar)
        out.append(f"update+finalise alone {t:.1f} us")
        cm.update(x); cm.update_finalise(x); torch.cuda.synchronize(); bar()
        t = timeit(lambda: A.mult(x, y), sync=bar)
        out.append(f"mult alone (both blocks) {t:.1f} us")
        idx = torch.from_numpy(cm.get("indexbuf").astype(np.int32)).to(dev)
        buf = torch.empty(len(idx), dtype=torch.float64, device=dev)
        t = timeit(lambda: K.gather(idx, x, buf))
        out.append(f"pack alone ({len(idx)} values) {t:.1f} us")
    for which, name in ((0, "local"), (1, "remote")):
        blk = A.local_block() if which == 0 else A.remote_block()
        if blk is None or len(blk["values"]) == 0:
            continue
        blk["nnz"] = len(blk["values"])
        rp = torch.from_numpy(blk["rowptr"]).to(dev)
        ci = torch.from_numpy(blk["colind"]).to(dev)
        v = torch.from_numpy(blk["values"]).to(dev)
        plan = K.CsrPlan(blk["rows"], blk["cols"], rp, False, nnz=blk["nnz"])
        info = plan.info()
        beta = 0.0 if which == 0 else 1.0
        t = timeit(lambda: K.spmv(plan, rp, ci, v, x, y, 1.0, beta))
        nb = 12 * blk["nnz"] + 4 * (rows + 1) + (16 if which == 0 else 16) * rows
        out.append(f"{name} block kernel alone: nnz {blk['nnz']} walk {info.get('walk')} tiles {info.get('num_tiles')} "
                   f"{t:.1f} us ({nb / t / 1e3:.0f} GB/s)")
        nonempty = int((np.diff(blk["rowptr"]) > 0).sum())
        out.append(f"{name} block rows with entries: {nonempty} of {rows}")
        plan.close()
    for r in range(size):
        if r == rank:
            print("\n".join(out), flush=True)
        bar()
    A.close()
    if size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
