#!/usr/bin/env python
"""halo_probe.py -- 2+ rank probe (torchrun): NCCL halo latency alone and under a resident SpMV kernel,
CUDA IPC peer mapping, copy-engine peer copy latency, stream memory ops.  Prints from rank 0."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from spmv_b200 import api  # noqa: E402


def ev():
    return torch.cuda.Event(enable_timing=True)


def main():
    rank, size, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", lr)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    api.init_world_from_torch(dist.new_group(backend="gloo"))
    ex = api.B200Executor.create(lr)
    ex.use_torch_stream()
    P = lambda *a: print(*a, flush=True) if rank == 0 else None  # noqa: E731

    for st, sym in ((27, False), (7, True)):
        A = api.create_laplace3d(ex, 256, 256, 256 * size, st, sym, api.P2P_NONBLOCKING)
        cm = A.col_map()
        rows = A.rows()
        x = torch.ones(rows + cm.num_ghosts(), dtype=torch.float64, device=dev)
        y = torch.empty(rows, dtype=torch.float64, device=dev)
        for _ in range(5):
            cm.update(x); A.mult(x, y)
        torch.cuda.synchronize(); dist.barrier()
        # (1) exchange alone
        e0, e1 = ev(), ev()
        e0.record()
        for _ in range(50):
            cm.update(x); cm.update_finalise(x)
        e1.record(); torch.cuda.synchronize()
        P(f"stencil {st} sym={sym}: update+finalise alone {e0.elapsed_time(e1) / 50 * 1e3:.1f} us")
        dist.barrier()
        # (2) step, and mult alone
        e0, e1 = ev(), ev()
        e0.record()
        for _ in range(50):
            cm.update(x); A.mult(x, y)
        e1.record(); torch.cuda.synchronize()
        t_step = e0.elapsed_time(e1) / 50
        dist.barrier()
        cm.update(x); torch.cuda.synchronize(); dist.barrier()
        e0, e1 = ev(), ev()
        e0.record()
        for _ in range(50):
            A.mult(x, y)   # no update in flight: the arrived event is already complete
        e1.record(); torch.cuda.synchronize()
        print(f"[rank {rank}] stencil {st} sym={sym}: update+mult {t_step * 1e3:.1f} us; mult alone "
              f"{e0.elapsed_time(e1) / 50 * 1e3:.1f} us", flush=True)
        dist.barrier()
        # (2b) as bench.py does it: an event pair around mult in every step
        kev = [(ev(), ev()) for _ in range(50)]
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = ev(), ev()
        e0.record()
        for a, b in kev:
            cm.update(x); a.record(); A.mult(x, y); b.record()
        e1.record(); torch.cuda.synchronize()
        ks = sorted(a.elapsed_time(b) for a, b in kev)
        print(f"[rank {rank}] stencil {st} sym={sym}: with per-step events: step {e0.elapsed_time(e1) / 50 * 1e3:.1f} us; "
              f"mult min/median/max {ks[0] * 1e3:.1f}/{ks[25] * 1e3:.1f}/{ks[-1] * 1e3:.1f} us", flush=True)
        dist.barrier()
        A.close()

    # (3) CUDA IPC + copy engine + stream memory ops
    try:
        from cuda import cuda as cu, cudart
        n = 65536
        buf = torch.zeros(2 * n + 64, dtype=torch.float64, device=dev)
        err, handle = cudart.cudaIpcGetMemHandle(buf.data_ptr())
        P("cudaIpcGetMemHandle:", err)
        handles = [None] * size
        dist.all_gather_object(handles, bytes(handle.reserved))
        peer = (rank + 1) % size
        h = cudart.cudaIpcMemHandle_t()
        h.reserved = handles[peer]
        err, pptr = cudart.cudaIpcOpenMemHandle(h, cudart.cudaIpcMemLazyEnablePeerAccess)
        P("cudaIpcOpenMemHandle:", err, hex(pptr) if pptr else None)
        # NB torch tensors sit at an offset inside a cudaMalloc'ed segment: the handle maps the segment base
        err, base, sz = cu.cuMemGetAddressRange(buf.data_ptr())
        off = buf.data_ptr() - int(base)
        offs = [None] * size
        dist.all_gather_object(offs, off)
        pdst = pptr + offs[peer]
        src = torch.full((n,), float(rank + 1), dtype=torch.float64, device=dev)
        s = torch.cuda.current_stream().cuda_stream
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = ev(), ev()
        e0.record()
        for _ in range(100):
            cudart.cudaMemcpyAsync(pdst, src.data_ptr(), n * 8, cudart.cudaMemcpyKind.cudaMemcpyDeviceToDevice, s)
        e1.record(); torch.cuda.synchronize(); dist.barrier()
        P(f"peer cudaMemcpyAsync 512 KiB via IPC mapping: {e0.elapsed_time(e1) / 100 * 1e3:.1f} us each; "
          f"landed={float(buf[0].item())} (expect {((rank - 1) % size) + 1})")
        # stream memory ops on the peer mapping: write a flag into the peer, wait for mine
        flag_off = 2 * n * 8
        err, = cu.cuStreamWriteValue32(s, pdst + flag_off, 7, 0)
        P("cuStreamWriteValue32 on peer mapping:", err)
        err, = cu.cuStreamWaitValue32(s, buf.data_ptr() + flag_off, 7, cu.CUstreamWaitValue_flags.CU_STREAM_WAIT_VALUE_GEQ)
        P("cuStreamWaitValue32:", err)
        t0 = time.time()
        torch.cuda.synchronize()
        P(f"flag wait completed in {time.time() - t0:.3f} s; flag word = {buf.view(torch.int32)[2 * n * 2].item()}")
        dist.barrier()
        cudart.cudaIpcCloseMemHandle(pptr)
    except Exception as e:  # noqa: BLE001
        import traceback
        traceback.print_exc()
        P("IPC probe failed:", repr(e))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
