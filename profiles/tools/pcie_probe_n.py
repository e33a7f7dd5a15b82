#!/usr/bin/env python
"""pcie_probe_n.py -- the host<->device ceiling that bounds `e2e` at N ranks (torchrun, one rank per GPU).

Every rank moves the e2e step's bytes (256^3 doubles up, 256^3 doubles down, pinned host memory) at the same
time as all the others: H2D alone, D2H alone, both at once.  The slowest rank's time for "both" is the floor of
Matrix::mult_host at that N (no kernel in it at all).  --bind pins each rank to the CPUs `nvidia-smi topo -m`
lists for its GPU before the pinned buffers are allocated (first touch on that socket).
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    bind = "--bind" in sys.argv
    rank, size, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    note = "unbound"
    if bind:
        import bench
        note = bench.bind_near_gpu(lr)
    dev = torch.device(f"cuda:{lr}")
    torch.cuda.set_device(dev)
    if size > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = 256 ** 3
    xh = torch.ones(n, dtype=torch.float64).pin_memory()
    yh = torch.empty(n, dtype=torch.float64).pin_memory()
    xd = torch.empty(n, dtype=torch.float64, device=dev)
    yd = torch.ones(n, dtype=torch.float64, device=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(h2d, d2h, reps=5):
        best = 1e9
        for _ in range(reps + 2):
            if size > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            s1.wait_stream(torch.cuda.current_stream())
            s2.wait_stream(torch.cuda.current_stream())
            if h2d:
                with torch.cuda.stream(s1):
                    xd.copy_(xh, non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    yh.copy_(yd, non_blocking=True)
            torch.cuda.current_stream().wait_stream(s1)
            torch.cuda.current_stream().wait_stream(s2)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            if size > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            best = min(best, float(t.item()))
        return best

    out = {"n_gpus": size, "bind": note if rank == 0 else None, "bytes_each_way_per_rank": 8 * n}
    for name, a, b in (("h2d_only", 1, 0), ("d2h_only", 0, 1), ("both", 1, 1)):
        ms = run(a, b)
        out[name] = {"ms_max_over_ranks": ms, "aggregate_GBps": (a + b) * 8 * n * size / ms / 1e6,
                     "per_rank_GBps_each_way": 8 * n / ms / 1e6}
    notes = [None] * size
    if size > 1:
        dist.all_gather_object(notes, note)
        dist.destroy_process_group()
    else:
        notes = [note]
    if rank == 0:
        out["bind_notes"] = notes
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
