#!/usr/bin/env python
"""bench.py -- the headline measurement (contract in the task prompt).

Workload (config.workload): distributed fp64 general-CSR SpMV, one halo update +
one Matrix::mult per step, on the 27-point 3-D Laplacian with 256x256x256 rows per
GPU (weak scaling: the grid is 256 x 256 x 256*N, 1-D slab partition).  That is the
case BASELINE.json's target is quoted on (>=70 % of the HBM roofline at 1 GPU).
metric = aggregate algorithmic GB/s (SURVEY.md section 8d formula).

Extra numbers ride in "also": symmetric-storage 7-point SpMV (config 2) and CG
iterations/s (configs 3/4 shape, 27-point 256^3 per GPU).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NX = NY = 256
NZ_PER_GPU = 256
STENCIL = 27


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nz", type=int, default=NZ_PER_GPU, help="z-planes per GPU (default 256)")
    ap.add_argument("--no-extras", action="store_true", help="skip the 'also' measurements and cpu_baseline")
    ap.add_argument("--sweep", action="store_true", help="kernel tuning sweep (prints a table, no JSON)")
    return ap.parse_args()


# ----------------------------------------------------------------------------
# clocks sampling (B200_PROFILING.md recipe)
# ----------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [t.strip() for t in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def algorithmic_bytes(nnz, rows, cols_touched):
    """General CSR SpMV, SURVEY.md 8d: 12*nnz + 4*(rows+1) + 8*cols_touched + 8*rows."""
    return 12 * nnz + 4 * (rows + 1) + 8 * cols_touched + 8 * rows


# ----------------------------------------------------------------------------
# the B200 arm
# ----------------------------------------------------------------------------
class SlabOperator:
    """One rank's share of the Laplacian on a 1-D slab partition, with the halo plan the
    reference's L2GMap would build for it (L2GMap.cpp:346-479): neighbours ascending by
    rank, ghosts landing in the tail of x, no unpack step (L2GMap.cpp:586-598)."""

    def __init__(self, rank, size, nx, ny, nz, stencil, dev, nccl_comm, symmetric=False):
        import torch
        from spmv_b200 import kernels as K

        self.K, self.torch = K, torch
        self.rank, self.size, self.dev, self.comm = rank, size, dev, nccl_comm
        z0, z1, nzg = rank * nz, (rank + 1) * nz, size * nz
        plane = nx * ny
        self.rows = plane * nz
        self.symmetric = symmetric
        mode = 1 if symmetric else (0 if size == 1 else 2)
        g = K.laplace3d(nx, ny, nzg, z0, z1, stencil, mode, dev)
        self.nghosts = g["nghosts"]
        self.ncols = self.rows + self.nghosts
        self.local, self.remote, self.diag = g["local"], g["remote"], g["diag"]
        self.nnz_local = int(self.local[0][-1].item())
        self.nnz_remote = int(self.remote[0][-1].item()) if self.remote is not None else 0
        self.plan_local = K.CsrPlan(self.rows, self.ncols, self.local[0], symmetric, nnz=self.nnz_local)
        self.plan_remote = (K.CsrPlan(self.rows, self.ncols, self.remote[0], False, nnz=self.nnz_remote)
                            if self.remote is not None and self.nnz_remote > 0 else None)
        # halo plan
        nbr, sidx, scnt, rcnt = [], [], [], []
        if rank > 0:
            nbr.append(rank - 1)
            sidx.append(torch.arange(0, plane, dtype=torch.int32))
            scnt.append(plane)
            rcnt.append(plane)
        if rank < size - 1:
            nbr.append(rank + 1)
            sidx.append(torch.arange(self.rows - plane, self.rows, dtype=torch.int32))
            scnt.append(plane)
            rcnt.append(plane)
        import numpy as np
        self.nnbr = len(nbr)
        self.nbr = np.array(nbr, dtype=np.int32)
        self.send_cnt = np.array(scnt, dtype=np.int32)
        self.send_off = np.concatenate([[0], np.cumsum(scnt)[:-1]]).astype(np.int32) if nbr else np.zeros(0, np.int32)
        self.recv_cnt = np.array(rcnt, dtype=np.int32)
        self.recv_off = (self.rows + np.concatenate([[0], np.cumsum(rcnt)[:-1]])).astype(np.int32) if nbr else np.zeros(0, np.int32)
        self.indexbuf = torch.cat(sidx).to(dev) if nbr else None
        self.sendbuf = torch.empty(int(sum(scnt)), dtype=torch.float64, device=dev) if nbr else None
        self.comm_stream = torch.cuda.Stream(device=dev) if nbr else None
        self.ev_packed = torch.cuda.Event()
        self.ev_arrived = torch.cuda.Event()
        self.launches_per_step = (1 if self.nnbr == 0 else 2) + (1 if self.plan_remote else 0) + (1 if symmetric else 0)

    def nnz_full(self):
        if self.symmetric:
            return 2 * self.nnz_local + self.rows + self.nnz_remote
        return self.nnz_local + self.nnz_remote

    def algorithmic_bytes(self):
        if self.symmetric:  # 12*nnz_lower + 4*(rows+1) + 3*8*rows (+ remote block)
            b = 12 * self.nnz_local + 4 * (self.rows + 1) + 24 * self.rows
            if self.nnz_remote:
                b += 12 * self.nnz_remote + 4 * (self.rows + 1) + 8 * self.nghosts
            return b
        b = algorithmic_bytes(self.nnz_local, self.rows, self.ncols)
        if self.nnz_remote:
            b += 12 * self.nnz_remote + 4 * (self.rows + 1)
        return b

    def update(self, x):
        """L2GMap::update (p2p_nonblocking flavour): pack, then grouped send/recv on a side stream."""
        if self.nnbr == 0:
            return
        import ctypes as C
        from spmv_b200 import check, lib
        torch, K = self.torch, self.K
        K.gather(self.indexbuf, x, self.sendbuf)
        main = torch.cuda.current_stream()
        self.ev_packed.record(main)
        self.comm_stream.wait_event(self.ev_packed)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        check(lib().b200_halo_exchange(self.comm, self.nnbr, p(self.nbr), C.c_void_p(self.sendbuf.data_ptr()),
                                       p(self.send_cnt), p(self.send_off), C.c_void_p(x.data_ptr()),
                                       p(self.recv_cnt), p(self.recv_off), 8,
                                       C.c_void_p(self.comm_stream.cuda_stream)))
        self.ev_arrived.record(self.comm_stream)

    def mult(self, x, y, dot=None, kernel_events=None):
        """Matrix::mult: spmv / spmv_overlap / spmv_sym(_overlap) schedules (Matrix.cpp:483-552)."""
        torch, K = self.torch, self.K
        if kernel_events is not None:
            kernel_events[0].record()
        if self.symmetric:
            K.symv(self.plan_local, *self.local, self.diag, x, y)
        elif dot is not None and self.plan_remote is None:
            K.spmv_dot(self.plan_local, *self.local, x, y, dot)
        else:
            K.spmv(self.plan_local, *self.local, x, y)
        if kernel_events is not None:
            kernel_events[1].record()
        if self.nnbr:
            torch.cuda.current_stream().wait_event(self.ev_arrived)  # update_finalise
        if self.plan_remote is not None:
            K.spmv(self.plan_remote, *self.remote, x, y, 1.0, 1.0)


def make_nccl_comm(rank, size, dev):
    """Our own NCCL communicator (ids travel over torch.distributed)."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    from spmv_b200 import check, lib
    if size == 1:
        return None
    buf = C.create_string_buffer(128)
    if rank == 0:
        check(lib().b200_nccl_unique_id(buf))
    t = torch.tensor(list(buf.raw), dtype=torch.uint8, device=dev)
    dist.broadcast(t, 0)
    raw = bytes(t.cpu().tolist())
    comm = C.c_void_p()
    check(lib().b200_nccl_comm_create(C.byref(comm), size, rank, raw))
    return comm


class Timer:
    """barrier + synchronize on both sides, CUDA events on the launching stream, max over ranks."""

    def __init__(self, size, dev):
        import torch
        self.torch, self.size, self.dev = torch, size, dev

    def run(self, fn, steps):
        torch = self.torch
        if self.size > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if self.size > 1:
            import torch.distributed as dist
            t = torch.tensor([ms], dtype=torch.float64, device=self.dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
            dist.barrier()
        return ms


def cg_iteration_factory(op, size, dev):
    """One CG iteration (cg.cpp:55-86) as fused kernels; returns (iterate, state, launches)."""
    import ctypes as C
    import torch
    from spmv_b200 import check, lib
    from spmv_b200 import kernels as K
    M = op.rows
    ones = torch.ones(op.ncols, dtype=torch.float64, device=dev)
    b = torch.empty(M, dtype=torch.float64, device=dev)
    op.update(ones)
    op.mult(ones, b)  # b = A*1
    p = torch.zeros(op.ncols, dtype=torch.float64, device=dev)
    p[:M] = b
    r, x, Ap = b.clone(), torch.zeros(M, dtype=torch.float64, device=dev), torch.empty(M, dtype=torch.float64, device=dev)
    st, scratch = K.new_cg_state(dev), K.new_dot_scratch(dev)
    rr = torch.zeros(1, dtype=torch.float64, device=dev)
    K.dot(r, r, rr, scratch)
    sp = lambda t: C.c_void_p(t.data_ptr())
    cur = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if size > 1:
        check(lib().b200_allreduce_sum_f64(op.comm, sp(rr), 1, cur()))
    K.cg_init_state(st, rr, 1e-30, 1 << 30)  # never converges inside the timed window
    fused = op.plan_remote is None and not op.symmetric
    pM = p[:M]

    def iterate(_i=0):
        op.update(p)
        if fused:
            op.mult(p, Ap, dot=st[1:2])
        else:
            op.mult(p, Ap)
            K.dot(pM, Ap, st[1:2], scratch, st[4:5])
        if size > 1:
            check(lib().b200_allreduce_sum_f64(op.comm, sp(st[1:2]), 1, cur()))
        K.cg_update_xr(M, st, pM, Ap, x, r)
        if size > 1:
            check(lib().b200_allreduce_sum_f64(op.comm, sp(st[2:3]), 1, cur()))
        K.cg_update_p(M, st, r, pM)

    launches = op.launches_per_step + (0 if fused else 1) + 2
    return iterate, st, launches


def run_b200(args):
    import numpy as np
    import torch
    from spmv_b200 import kernels as K
    from spmv_b200 import synth

    rank = int(os.environ.get("RANK", "0"))
    size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert size == args.gpus or size == 1 and args.gpus == 1, f"--gpus {args.gpus} but WORLD_SIZE={size}"
    assert torch.cuda.is_available(), "bench.py needs a GPU: there is no CPU path"
    dev = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    if size > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    comm = make_nccl_comm(rank, size, dev)
    timer = Timer(size, dev)
    nz = args.nz

    op = SlabOperator(rank, size, NX, NY, nz, STENCIL, dev, comm)
    if args.sweep:
        return sweep(op, dev, args)
    M, N_glob = op.rows, op.rows * size
    x = torch.zeros(op.ncols, dtype=torch.float64, device=dev)
    x[:M] = torch.from_numpy(synth.bump_vector(N_glob, rank * M, (rank + 1) * M)).to(dev)
    y = torch.empty(M, dtype=torch.float64, device=dev)

    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]

    def step(i):
        op.update(x)
        op.mult(x, y, kernel_events=kev[i] if i < len(kev) else None)

    for i in range(max(args.warmup, 3)):
        step(10 ** 9)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    ms = timer.run(step, args.steps)
    # keep the GPU busy a little longer so the 100 ms sampler sees the loaded clocks
    if rank == 0:
        t_end = time.time() + 1.0
        while time.time() < t_end:
            for i in range(20):
                step(10 ** 9)
            torch.cuda.synchronize()
        clocks = sampler.stop()
    ms_step = ms / args.steps
    bytes_rank = op.algorithmic_bytes()
    value = bytes_rank * size / (ms_step * 1e-3) / 1e9
    kern_ms = sorted(a.elapsed_time(b) for a, b in kev)
    kern_avg = sum(kern_ms) / len(kern_ms)
    kern_bytes = algorithmic_bytes(op.nnz_local, op.rows, op.rows if size > 1 else op.ncols)
    peak, peak_src = peak_hbm()
    achieved = kern_bytes / (kern_avg * 1e-3) / 1e9

    # ---- e2e: host buffers, H2D of x and D2H of y inside the timed region -------------
    xh = torch.empty(M, dtype=torch.float64).pin_memory()
    xh.copy_(x[:M].cpu())
    yh = torch.empty(M, dtype=torch.float64).pin_memory()

    def e2e_step(i):
        x[:M].copy_(xh, non_blocking=True)
        op.update(x)
        op.mult(x, y)
        yh.copy_(y, non_blocking=True)

    e2e_steps = max(3, min(args.steps, 10))
    for i in range(3):
        e2e_step(i)
    ms_e2e = timer.run(e2e_step, e2e_steps) / e2e_steps
    e2e = {"value": bytes_rank * size / (ms_e2e * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": ms_e2e,
           "h2d_bytes_per_step": 8 * M, "d2h_bytes_per_step": 8 * M}

    also = {}
    if not args.no_extras:
        # CG iterations/s on the same operator (configs 3/4 shape)
        iterate, st, cg_launches = cg_iteration_factory(op, size, dev)
        for i in range(3):
            iterate()
        cg_steps = 20
        ms_cg = timer.run(iterate, cg_steps) / cg_steps
        also["cg_27pt"] = {"iters_per_s": 1e3 / ms_cg, "ms_per_iter": ms_cg, "rows_per_gpu": M,
                           "gpu_launches_per_iter": cg_launches,
                           "algorithmic_GBps": (bytes_rank + 11 * 8 * M) * size / (ms_cg * 1e-3) / 1e9}
        del iterate, st
        # symmetric-storage 7-point SpMV (config 2)
        ops = SlabOperator(rank, size, NX, NY, nz, 7, dev, comm, symmetric=True)
        ys = torch.empty(M, dtype=torch.float64, device=dev)

        def sstep(i):
            ops.update(x)
            ops.mult(x, ys)

        for i in range(3):
            sstep(i)
        ms_s = timer.run(sstep, 20) / 20
        sb = ops.algorithmic_bytes()
        also["sym_7pt"] = {"GBps": sb * size / (ms_s * 1e-3) / 1e9, "ms_per_step": ms_s,
                           "GFLOPs": 2 * ops.nnz_full() * size / (ms_s * 1e-3) / 1e9,
                           "frac_of_hbm_peak": sb / (ms_s * 1e-3) / 1e9 / peak}
        del ops

    if rank != 0:
        return
    line = {
        "metric": "spmv_algorithmic_GBps", "value": value, "unit": "GB/s", "n_gpus": size, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"general CSR SpMV (halo update + Matrix::mult), 27-point Laplacian "
                               f"{NX}x{NY}x{nz} rows per GPU, 1-D slab partition",
                   "rows_per_gpu": M, "nnz_per_gpu": op.nnz_local + op.nnz_remote, "n_gpus": size,
                   "l2": "inputs (5.7 GB matrix stream per step) far exceed the 126 MB L2; no explicit flush",
                   "GFLOPs": 2 * (op.nnz_local + op.nnz_remote) * size / (ms_step * 1e-3) / 1e9},
        "e2e": e2e,
        "gpu_launches": op.launches_per_step * args.steps,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "kernel": "spmv_pipe_kernel<double> (local block)",
                     "kernel_ms_avg": kern_avg, "kernel_ms_min": kern_ms[0], "algorithmic_bytes": kern_bytes,
                     "peak_source": peak_src, "frac_of_8TBps_nominal": achieved / 8000.0},
        "also": also,
    }
    if size == 1 and not args.no_extras:
        line["cpu_baseline"] = cpu_baseline(sample_nz=32)
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------
# CPU legs: the reference's own OpenMP kernel (oracle/_ref), all host threads
# ----------------------------------------------------------------------------
def _cpu_operator(sample_nz):
    import numpy as np
    from oracle import kernels as OK
    from spmv_b200 import synth
    rp, col, val = synth.laplace3d_rows(NX, NY, sample_nz, 0, sample_nz, STENCIL)
    rows = len(rp) - 1
    cores = os.cpu_count() or 1
    if OK.have_ref():
        op = OK.RefOmp(rows, rows, rp, col, val, threads=cores)
        kind = "reference"
        run = lambda x, y: op.run(x, y)
    else:  # the plain-C port, one thread
        rp32, c32 = rp.astype(np.int32), col.astype(np.int32)
        kind, cores = "port", 1
        run = lambda x, y: OK.lib().oracle_csr_spmv_f64(rows, rp32, c32, val, 1.0, x, 0.0, y)
    x = synth.bump_vector(rows, 0, rows)
    y = np.zeros(rows)
    nbytes = algorithmic_bytes(int(rp[-1]), rows, rows)
    return run, x, y, nbytes, kind, cores, rows


def cpu_baseline(sample_nz=32, budget_s=10.0):
    run, x, y, nbytes, kind, cores, rows = _cpu_operator(sample_nz)
    run(x, y)
    t0, n = time.perf_counter(), 0
    while True:
        run(x, y)
        n += 1
        dt = time.perf_counter() - t0
        if dt > budget_s or n >= 200:
            break
    return {"value": nbytes * n / dt / 1e9, "unit": "GB/s", "cores": cores, "kind": kind,
            "ms_per_spmv": dt / n * 1e3,
            "sample": f"27-point Laplacian {NX}x{NY}x{sample_nz} ({rows} rows, 1/{NZ_PER_GPU // sample_nz} of the "
                      f"1-GPU workload), general CSR, {n} SpMVs, "
                      + ("reference openmp/csr_kernels.openmp.cpp" if kind == "reference" else "oracle C port")}


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    sample_nz = 32
    run, x, y, nbytes, kind, cores, rows = _cpu_operator(sample_nz)
    for _ in range(max(args.warmup, 1)):
        run(x, y)
    steps = min(args.steps, 50)
    t0 = time.perf_counter()
    for _ in range(steps):
        run(x, y)
    dt = time.perf_counter() - t0
    v = nbytes * steps / dt / 1e9
    sample = (f"each step = one SpMV on a 27-point Laplacian {NX}x{NY}x{sample_nz} slab ({rows} rows; the GPU arm "
              f"runs {NZ_PER_GPU} planes per GPU), GB/s is size-independent; "
              + ("reference openmp/csr_kernels.openmp.cpp, OMP_NUM_THREADS=%d" % cores if kind == "reference"
                 else "oracle C port, 1 thread"))
    print(json.dumps({
        "impl": "reference", "metric": "spmv_algorithmic_GBps", "value": v, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": max(args.warmup, 1), "ms_per_step": dt / steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "general CSR SpMV, 27-point Laplacian, CPU sample", "rows": rows},
        "cpu_baseline": {"value": v, "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


# ----------------------------------------------------------------------------
# tuning sweep (one gpurun call explores the kernel's launch shape)
# ----------------------------------------------------------------------------
def sweep(op, dev, args):
    import itertools
    import torch
    from spmv_b200 import synth
    M = op.rows
    x = torch.zeros(op.ncols, dtype=torch.float64, device=dev)
    x[:M] = torch.from_numpy(synth.bump_vector(M, 0, M)).to(dev)
    y = torch.empty(M, dtype=torch.float64, device=dev)
    nbytes = op.algorithmic_bytes()
    peak, _ = peak_hbm()
    rows = []
    grid = [(0, tb, st, cw, ct) for tb in (16384, 24576, 32768, 49152, 65536) for st in (2, 3, 4)
            for cw in (2, 4, 8) for ct in (1, 2, 3, 4)]
    grid += [(1, tb, 2, 4, 1) for tb in (8192, 16384, 32768, 49152)]
    for variant, tb, st, cw, ct in grid:
        smem = st * (tb + 2048 + 1056) + 16 * st + 256
        if variant == 0 and smem * ct > 227 * 1024:
            continue
        try:
            op.plan_local.tune(variant, tb, st, cw, ct)
            for _ in range(3):
                op.mult(x, y)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                op.mult(x, y)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
        except Exception as e:  # noqa: BLE001
            print(f"variant={variant} tile={tb} stages={st} cw={cw} ctas={ct}: {e}", flush=True)
            continue
        rows.append((ms, variant, tb, st, cw, ct))
        print(f"variant={variant} tile={tb:6d} stages={st} cw={cw} ctas={ct}: {ms:.4f} ms "
              f"{nbytes / ms / 1e6:8.1f} GB/s  {nbytes / ms / 1e6 / peak:.3f} of peak", flush=True)
    rows.sort()
    print("BEST:", rows[:5], flush=True)


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    run_b200(args)
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass


if __name__ == "__main__":
    main()
