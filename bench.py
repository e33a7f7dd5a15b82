#!/usr/bin/env python
"""bench.py -- the headline measurement (contract in the task statement).

Workload (config.workload) = BASELINE.json configs[1]: symmetric-storage fp64 CSR SpMV on
the 3-D 7-point Laplacian, 256^3 rows per GPU (weak scaling over z-slabs: the grid is
256 x 256 x 256*N).  One step = L2GMap::update + Matrix::mult, called through the
reference-shaped API (spmv_b200.api -> libspmv_b200.so -> the C ABI of libb200spmv.so).
metric = aggregate algorithmic GB/s (SURVEY.md 8d byte formulas).

"also" carries the other BASELINE configs that fit this run: general 27-point SpMV (the
case the >=70 % roofline target is quoted on), CG iterations/s and the CG solve to
rtol 1e-8 on the 27-point operator (configs 3/4 shape), each with its own roofline figure.
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
CONFIG5_ROWS_PER_GPU = 6_250_000  # BASELINE configs[4]: ~50 M rows on 8 GPUs


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nz", type=int, default=NZ_PER_GPU, help="z-planes per GPU (default 256)")
    ap.add_argument("--strong", action="store_true",
                    help="strong scaling (BASELINE configs[2]): the 256^3 grid split over the GPUs, nz = 256 / N")
    ap.add_argument("--no-extras", action="store_true", help="skip the 'also' measurements and cpu_baseline")
    ap.add_argument("--no-library-context", action="store_true",
                    help="N=1: skip the cuSPARSE SpMV / cuBLAS BLAS-1 context timings (what the reference's CUDA backend "
                         "runs, cuda/csr_kernels.cuda.cu:106-151, cuda/cg.cuda.cu:88-150); library code, context only")
    ap.add_argument("--config5-rows", type=int, default=CONFIG5_ROWS_PER_GPU,
                    help="rows per GPU of the power-law / PETSc leg (default 6.25 M = 50 M rows at 8 GPUs; 0 skips it)")
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


# ----------------------------------------------------------------------------
# algorithmic bytes (x counted once, y written once)
# ----------------------------------------------------------------------------
def bytes_general(nnz_local, nnz_remote, rows, ncols, two_blocks):
    b = 12 * nnz_local + 4 * (rows + 1) + 8 * ncols + 8 * rows
    if two_blocks:
        b += 12 * nnz_remote + 4 * (rows + 1)
    return b


def bytes_symmetric(nnz_lower, nnz_remote, rows, nghosts):
    b = 12 * nnz_lower + 4 * (rows + 1) + 24 * rows
    if nnz_remote:
        b += 12 * nnz_remote + 4 * (rows + 1) + 8 * nghosts
    return b


# ----------------------------------------------------------------------------
# the B200 arm: everything through the reference-shaped API (spmv_b200.api)
# ----------------------------------------------------------------------------
class Operator:
    """One rank's spmv::Matrix of the slab-partitioned Laplacian plus its vectors."""

    def __init__(self, api, ex, dev, rank, size, nz, stencil, symmetric, matrix=None, row0=None, n_glob=None):
        import torch
        from spmv_b200 import synth
        self.api, self.ex, self.torch = api, ex, torch
        self.symmetric, self.stencil, self.size = symmetric, stencil, size
        # overlap schedule (local block while the halo is in flight) whenever there are neighbours
        cm = api.P2P_NONBLOCKING if size > 1 else api.P2P_BLOCKING
        self.A = matrix if matrix is not None else api.create_laplace3d(ex, NX, NY, nz * size, stencil, symmetric, cm)
        info = self.A._info()
        self.rows, self.nnz_local, self.nnz_remote = info[0], info[6], info[7]
        self.two_blocks = bool(info[5])
        self.col_map = self.A.col_map()
        self.nghosts = self.col_map.num_ghosts()
        self.ncols = self.rows + self.nghosts
        self.nnbr = len(self.col_map.get("neighbours")) if size > 1 else 0
        n_glob = self.rows * size if n_glob is None else n_glob
        row0 = rank * self.rows if row0 is None else row0
        self.x = torch.zeros(self.ncols, dtype=torch.float64, device=dev)
        self.x[: self.rows] = torch.from_numpy(synth.bump_vector(n_glob, row0, row0 + self.rows)).to(dev)
        self.y = torch.empty(self.rows, dtype=torch.float64, device=dev)
        # our kernels per step: pack (if neighbours) + local block (symmetric storage: + the zero-fill of y,
        # zero_fill_desc_kernel) + remote block (if it has entries)
        self.launches_per_step = ((1 if self.nnbr else 0) + 1 + (1 if symmetric else 0)
                                  + (1 if self.nnz_remote > 0 else 0))

    def nnz_full(self):
        if self.symmetric:
            return 2 * self.nnz_local + self.rows + self.nnz_remote
        return self.nnz_local + self.nnz_remote

    def algorithmic_bytes(self):
        if self.symmetric:
            return bytes_symmetric(self.nnz_local, self.nnz_remote, self.rows, self.nghosts)
        return bytes_general(self.nnz_local, self.nnz_remote, self.rows, self.ncols, self.two_blocks)

    def local_kernel_bytes(self):
        """Algorithmic bytes of the dominant launch (the local block)."""
        if self.symmetric:
            return bytes_symmetric(self.nnz_local, 0, self.rows, 0)
        return bytes_general(self.nnz_local, 0, self.rows, self.rows if self.two_blocks else self.ncols, False)

    def step(self, kernel_events=None):
        self.col_map.update(self.x)  # L2GMap::update (demos/spmv.cpp:88)
        if kernel_events is not None:
            kernel_events[0].record()
        self.A.mult(self.x, self.y)  # Matrix::mult   (demos/spmv.cpp:93)
        if kernel_events is not None:
            kernel_events[1].record()

    def close(self):
        self.A.close()


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


def measure_spmv(op, timer, steps, warmup, peak, with_e2e):
    """Timed loop of update+mult; returns dict with step time, kernel time, optional e2e."""
    import torch
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    # >= 10 untimed steps: the first exchanges also set up NCCL's peer connections (milliseconds)
    for _ in range(max(warmup, 10 if op.size > 1 else 3)):
        op.step()
    ms = timer.run(lambda i: op.step(kev[i]), steps)
    ms_step = ms / steps
    kern = sorted(a.elapsed_time(b) for a, b in kev)
    kern_avg = sum(kern) / len(kern)
    out = {"ms_per_step": ms_step, "kernel_ms_avg": kern_avg, "kernel_ms_min": kern[0],
           "kernel_ms_median": kern[len(kern) // 2],
           "GBps": op.algorithmic_bytes() * op.size / (ms_step * 1e-3) / 1e9,
           "GFLOPs": 2 * op.nnz_full() * op.size / (ms_step * 1e-3) / 1e9,
           "kernel_GBps": op.local_kernel_bytes() / (kern_avg * 1e-3) / 1e9, "local_kernel_bytes": op.local_kernel_bytes()}
    out["frac_of_hbm_peak"] = out["kernel_GBps"] / peak
    if with_e2e:
        # host buffers: H2D of x and D2H of y inside the timed region.  Matrix::mult_host is
        # copy_from + update + mult + copy_to (demos/spmv_cuda.cpp:71-110) in one call, with the
        # copies pipelined against the tile ranges when the rank has no neighbours
        M = op.rows
        xh = torch.empty(M, dtype=torch.float64).pin_memory()
        xh.copy_(op.x[:M].cpu())
        yh = torch.empty(M, dtype=torch.float64).pin_memory()
        xh_np, yh_np = xh.numpy(), yh.numpy()

        def e2e_step(_i):
            op.A.mult_host(xh_np, yh_np)

        n = max(3, min(steps, 10))
        for i in range(3):
            e2e_step(i)
        ms_e2e = timer.run(e2e_step, n) / n
        op.step()
        err = float((torch.from_numpy(yh_np) - op.y.cpu()).abs().max() / op.y.abs().max().cpu())
        if not err <= 1e-13:
            raise RuntimeError(f"mult_host disagrees with update+mult: rel max err {err}")
        # the link's floor for this step: the same bytes up and down at once on two streams, every rank at the same
        # time, no kernel at all (what profiles/tools/pcie_probe_n.py measures stand-alone)
        s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
        xd, yd = op.x[:M], op.y

        def floor_step(_i):
            cur = torch.cuda.current_stream()
            s1.wait_stream(cur)
            s2.wait_stream(cur)
            with torch.cuda.stream(s1):
                xd.copy_(xh, non_blocking=True)
            with torch.cuda.stream(s2):
                yh.copy_(yd, non_blocking=True)
            cur.wait_stream(s1)
            cur.wait_stream(s2)

        for i in range(2):
            floor_step(i)
        ms_floor = timer.run(floor_step, 5) / 5
        out["e2e"] = {"value": op.algorithmic_bytes() * op.size / (ms_e2e * 1e-3) / 1e9, "unit": "GB/s",
                      "ms_per_step": ms_e2e, "h2d_bytes_per_step": 8 * M, "d2h_bytes_per_step": 8 * M,
                      "pcie_floor_ms": ms_floor, "over_floor": ms_e2e / ms_floor,
                      "floor_how": "H2D of x and D2H of y (the step's bytes) concurrently on two streams, all ranks at once, "
                                   "no kernel; max over ranks"}
    return out


def laplace_parity(op, rank, size, nz, stencil, dev, planes=4):
    """In-run parity of a Laplacian leg at bench size and at every N: regenerate sampled z-planes of THIS
    rank's rows on the host (synth.laplace3d_rows gives any plane range with global columns; x is a formula
    of the global index, so ghost values are known too), sum each row left to right in fp64
    (csr_kernels.cpp:42-49) and compare with the y the timed loop left on the device.  The first and last
    plane of the slab are always in the sample: those are the rows that read ghosts.  Raises above 1e-12
    (BASELINE tolerance: relative L2)."""
    import numpy as np
    import torch
    from spmv_b200 import synth
    plane = NX * NY
    nzg = nz * size
    n_glob = plane * nzg
    z0 = rank * nz
    rng = np.random.default_rng(20261105 + rank)
    zs = sorted({z0, z0 + nz - 1, z0 + nz // 2, z0 + int(rng.integers(0, nz))})[:max(planes, 2)]
    op.step()  # y = A x for the x of this operator
    y = op.y
    num = den = 0.0
    worst_row = 0.0
    for z in zs:
        rp, col, val = synth.laplace3d_rows(NX, NY, nzg, z, z + 1, stencil)
        xs = np.exp(-10.0 * np.power(5.0 * (col.astype(np.float64) / float(n_glob) - 0.5), 2.0))
        prod = val * xs
        want = np.zeros(plane)
        cnt = np.diff(rp)
        for j in range(int(cnt.max())):  # left to right, one entry of every row at a time
            m = cnt > j
            want[m] += prod[rp[:-1][m] + j]
        got = y[(z - z0) * plane:(z - z0 + 1) * plane].cpu().numpy()
        num += float(np.sum((got - want) ** 2))
        den += float(np.sum(want ** 2))
        scale = np.maximum(np.abs(want), np.abs(want).max() * 1e-30 + 1e-300)
        worst_row = max(worst_row, float(np.max(np.abs(got - want) / scale)))
    t = torch.tensor([num, den], dtype=torch.float64, device=dev)
    w = torch.tensor([worst_row], dtype=torch.float64, device=dev)
    if size > 1:
        import torch.distributed as dist
        dist.all_reduce(t)
        dist.all_reduce(w, op=dist.ReduceOp.MAX)
    rel = float(torch.sqrt(t[0] / t[1]).item()) if float(t[1].item()) > 0 else float("nan")
    out = {"rel_l2": rel, "max_rel_err_row": float(w.item()), "rows_checked_per_rank": len(zs) * plane,
           "planes_per_rank": [int(z - z0) for z in zs], "tolerance": 1e-12,
           "how": "sampled z-planes (first, last, middle, one random) of every rank regenerated on the host, "
                  "left-to-right fp64 row sums, relative L2 over all ranks' samples"}
    if not rel <= 1e-12:
        raise RuntimeError(f"{stencil}-point Laplacian: y differs from the regenerated rows, rel L2 {rel}")
    return out


def measure_config5(api, ex, dev, rank, size, timer, peak, rows_per_gpu):
    """BASELINE configs[4]: irregular power-law rows, written as a PETSc binary file by the host-core
    generator (demos/CreatePowerLaw.h), read back with read_petsc_binary_matrix, general CSR SpMV
    (update + mult).  6.25 M rows per GPU: the 50 M-row matrix at 8 GPUs."""
    import numpy as np
    import shutil
    import tempfile
    import torch
    from spmv_b200 import synth
    n = rows_per_gpu * size
    root = "/dev/shm" if os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > 110 * n + (1 << 30) \
        else tempfile.gettempdir()
    path = os.path.join(root, f"spmv_b200_config5_{os.environ.get('MASTER_PORT', str(os.getpid()))}.petsc")
    t_gen = nnz_glob = file_bytes = None
    if rank == 0:
        t0 = time.perf_counter()
        nnz_glob = api.write_powerlaw_petsc(path, n)
        t_gen = time.perf_counter() - t0
    if size > 1:
        import torch.distributed as dist
        dist.barrier()
    try:
        t0 = time.perf_counter()
        cm = api.P2P_NONBLOCKING if size > 1 else api.P2P_BLOCKING
        A = api.read_petsc_binary_matrix(path, ex, False, cm)
        torch.cuda.synchronize()
        t_read = time.perf_counter() - t0
        if size > 1:
            import torch.distributed as dist
            dist.barrier()
    finally:
        if rank == 0 and os.path.exists(path):
            file_bytes = os.path.getsize(path)
            os.remove(path)
    r0 = int(synth.owner_ranges(size, n)[rank])
    op = Operator(api, ex, dev, rank, size, 0, 0, False, matrix=A, row0=r0, n_glob=n)
    m = measure_spmv(op, timer, 30, 5, peak, with_e2e=False)
    # parity at full size: any rank can regenerate any row, and x is a formula of the global index --
    # recompute a sample of this rank's rows on the host (left-to-right sums, csr_kernels.cpp:42-49)
    rng = np.random.default_rng(20261104 + rank)
    y = op.y.cpu().numpy()
    worst = 0.0
    picks = np.unique(np.concatenate([rng.integers(0, op.rows, 256), [0, op.rows - 1]]))
    for lr in picks:
        rp, col, val = api.powerlaw_rows(n, r0 + int(lr), r0 + int(lr) + 1)
        xs = np.exp(-10.0 * np.power(5.0 * (col.astype(np.float64) / float(n) - 0.5), 2.0))
        want = 0.0
        for v, xv in zip(val, xs):
            want += v * xv
        worst = max(worst, abs(want - y[lr]) / max(abs(want), 1e-300))
    if not worst <= 1e-12:
        raise RuntimeError(f"config 5: y differs from the regenerated rows, rel err {worst}")
    t = torch.tensor([float(op.nghosts), float(op.nnbr), worst, t_read], dtype=torch.float64, device=dev)
    if size > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out = {"rows_global": n, "rows_per_gpu": op.rows, "nnz_global": nnz_glob, "GBps": m["GBps"], "GFLOPs": m["GFLOPs"],
           "ms_per_step": m["ms_per_step"], "local_kernel_ms_avg": m["kernel_ms_avg"],
           "local_kernel_GBps": m["kernel_GBps"], "local_kernel_frac_of_hbm_peak": m["frac_of_hbm_peak"],
           "algorithmic_bytes_rank0": op.algorithmic_bytes(), "max_ghosts_per_rank": int(t[0].item()),
           "max_neighbours": int(t[1].item()), "halo_bytes_in_per_rank_max": 8 * int(t[0].item()),
           "generate_and_write_s": t_gen, "read_petsc_s_max": float(t[3].item()),
           "file_bytes": file_bytes if rank == 0 else None,
           "parity": {"rows_checked_per_rank": int(len(picks)), "max_rel_err": float(t[2].item()),
                      "how": "sampled rows regenerated on the host, sequential fp64 row sums, tolerance 1e-12"},
           "input": "synth.powerlaw_rows definition: Pareto(2, 4) row lengths clamped to [1, 4096], 90 % of the "
                    "columns in a +-32768 band, 10 % anywhere; PETSc binary file -> read_petsc_binary_matrix"}
    op.close()
    return out


def measure_config1(api, ex, dev, rank, size, timer, n=1_000_000, n_apply=100):
    """BASELINE configs[0] as the reference's driver runs it (demos/spmv.cpp:73-108) on create_A(N)
    (demos/CreateA.cpp:31-103): warm-up, then n_apply x {L2GMap::update; Matrix::mult; x <- y}, final ||x||_2.
    24 MB of matrix: a latency-bound case -- what is reported is us per step and what it is made of.
    Parity: the same 100 products recomputed on the host in numpy, row sums left to right (the operator is
    tridiagonal: three vector expressions); bit-identical at one rank, <= 1e-12 otherwise (a rank's ghost column
    is numbered last, so its boundary row adds in another order)."""
    import numpy as np
    import torch
    from spmv_b200 import synth
    A = api.create_A(n, ex)
    cm = A.col_map()
    M, r0 = A.rows(), cm.global_offset()
    x = torch.zeros(M + cm.num_ghosts(), dtype=torch.float64, device=dev)
    x[:M] = torch.from_numpy(synth.bump_vector(n, r0, r0 + M)).to(dev)
    y = torch.empty(M, dtype=torch.float64, device=dev)
    cm.update(x)  # warm-up (demos/spmv.cpp:81-82)
    A.mult(x, y)
    y_first = y.cpu().numpy()

    def step(_i):
        cm.update(x)
        A.mult(x, y)
        x[:M].copy_(y)

    ms = timer.run(step, n_apply)
    g = 0.1

    def apply(v):  # rows [g, 1-2g, g], first [1-g, g], last [g, 1-g]; sums left to right like csr_kernels.cpp:45-47
        out = np.empty_like(v)
        out[1:-1] = (g * v[:-2] + (1.0 - 2.0 * g) * v[1:-1]) + g * v[2:]
        out[0] = (1.0 - g) * v[0] + g * v[1]
        out[-1] = g * v[-2] + (1.0 - g) * v[-1]
        return out

    xg = synth.bump_vector(n, 0, n)
    want_first = apply(xg)[r0:r0 + M]
    for _ in range(n_apply):
        xg = apply(xg)
    got = x[:M].cpu().numpy()
    want = xg[r0:r0 + M]
    e = torch.tensor([float(np.sum((got - want) ** 2)), float(np.sum(want ** 2)), float(np.sum(got ** 2)),
                      float(np.sum((y_first - want_first) ** 2)), float(np.sum(want_first ** 2))], dtype=torch.float64, device=dev)
    if size > 1:
        import torch.distributed as dist
        dist.all_reduce(e)
    rel = float(torch.sqrt(e[0] / e[1]).item())
    rel_first = float(torch.sqrt(e[3] / e[4]).item())
    if not (rel <= 1e-12 and rel_first <= 1e-12):
        raise RuntimeError(f"config 1: x after {n_apply} products differs from the host recomputation ({rel}, first y {rel_first})")
    info = A._info()
    nbytes = bytes_general(n * 3 - 2, 0, n, n, False)
    out = {"rows_global": n, "n_apply": n_apply, "us_per_step": ms / n_apply * 1e3, "GBps": nbytes / (ms / n_apply * 1e-3) / 1e9,
           "final_x_norm": float(torch.sqrt(e[2]).item()), "final_x_norm_host": float(np.linalg.norm(xg)),
           "parity": {"rel_l2_final_x": rel, "rel_l2_first_y": rel_first,
                      "bit_identical": bool(size == 1 and np.array_equal(got, want) and np.array_equal(y_first, want_first)),
                      "tolerance": 1e-12, "how": "100 products recomputed on the host, left-to-right row sums"},
           "halo_mode": cm.halo_mode(), "launches_per_step": 2 + (2 if size > 1 else 0),
           "bound": "latency: 12 MB of matrix per GPU stream in ~3 us; a step is 2-4 kernel launches + one device copy "
                    "(+ the neighbour's flag over NVLink), not bandwidth",
           "local_nnz": int(info[6])}
    A.close()
    return out


def measure_cg(api, op, timer, dev, size):
    """CG on `op` (b = A*1, x0 = 0): iterations/s over a fixed 100 iterations, then the solve
    to rtol 1e-8 (BASELINE config 3: kmax 5000) with its iteration count and final residual."""
    import torch
    M = op.rows
    ones = torch.zeros(op.ncols, dtype=torch.float64, device=dev)
    ones[:M] = 1.0
    b = torch.empty(M, dtype=torch.float64, device=dev)
    op.col_map.update(ones)
    op.A.mult(ones, b)
    xs = torch.zeros(M, dtype=torch.float64, device=dev)
    api.cg(op.ex, op.A, b, xs, 10, 0.0)  # warm-up
    iters = 100
    ms = timer.run(lambda i: api.cg(op.ex, op.A, b, xs, iters, 0.0), 1)
    ms2 = timer.run(lambda i: api.cg(op.ex, op.A, b, xs, iters, 0.0), 1)  # second call: the figure must reproduce
    # per-phase device time (profile mode: no graph, one iteration at a time, CUDA events between the parts)
    _, ph = api.cg_phases(op.ex, op.A, b, xs, 20, 0.0)
    pht = torch.tensor([ph["update"], ph["local"], ph["remote"], ph["vector"]], dtype=torch.float64, device=dev)
    if size > 1:
        import torch.distributed as dist
        dist.all_reduce(pht, op=dist.ReduceOp.MAX)
    ph = dict(zip(("halo_update_us", "local_spmv_us", "halo_wait_and_remote_spmv_us", "vector_step_and_allreduces_us"),
                  [float(v) for v in pht.tolist()]))
    ph["sum_us"] = sum(ph.values())
    ph["launch_gaps_and_skew_us"] = min(ms, ms2) / iters * 1e3 - ph["sum_us"]
    ph["how"] = ("spmv::cg profile mode, max over ranks of the per-iteration mean; gaps = timed ms_per_iter (graph path) "
                 "minus the sum: negative means the graph / overlap hides more than the profile mode's serial launches cost")
    res = {"iters_per_s": iters / (min(ms, ms2) * 1e-3), "ms_per_iter": min(ms, ms2) / iters, "rows_per_gpu": M,
           "ms_per_iter_runs": [ms / iters, ms2 / iters], "halo_mode": op.col_map.halo_mode(), "phases": ph,
           "algorithmic_GBps": (op.algorithmic_bytes() + 11 * 8 * M) * size / (ms / iters * 1e-3) / 1e9,
           "note": "100 iterations through spmv::cg incl. its setup (r0 norm; the workspace and the captured graph are "
                   "cached on the executor after the warm-up call), best of two calls"}
    holder = {}

    def solve(_i):
        holder["it"], holder["hist"] = api.cg(op.ex, op.A, b, xs, 5000, 1e-8, history=True)

    ms = timer.run(solve, 1)
    hist = holder["hist"]

    def gsum(v):  # sum over the ranks of a device scalar
        if size > 1:
            import torch.distributed as dist
            dist.all_reduce(v)
        return float(v.item())

    bnorm = gsum(torch.dot(b, b)) ** 0.5
    # ||b - A x|| / ||b|| recomputed with one more SpMV: the recurrence's residual can drift from the true one
    xp = torch.zeros(op.ncols, dtype=torch.float64, device=dev)
    xp[:M] = xs
    ax = torch.empty(M, dtype=torch.float64, device=dev)
    op.col_map.update(xp)
    op.A.mult(xp, ax)
    rtrue = gsum(torch.dot(b - ax, b - ax)) ** 0.5 / bnorm
    res["solve_rtol_1e-8"] = {"iterations": int(holder["it"]), "ms": ms,
                              "iters_per_s": holder["it"] / (ms * 1e-3) if ms > 0 else None,
                              "final_rel_residual": float(hist[-1] / bnorm) if len(hist) else None,
                              "residual_true": rtrue, "converged": bool(len(hist) and hist[-1] / bnorm < 1e-8)}
    if not rtrue < 1e-6:
        raise RuntimeError(f"CG: true residual {rtrue} after {holder['it']} iterations")
    if getattr(op, "symmetric", False) and size == 1:
        # x'Ax in the symmetric SpMV's epilogue (default) against the separate dot kernel: speed and how far the
        # residual histories part (they differ by roundings; so do two runs of symmetric storage with each other)
        prev = os.environ.get("B200_CG_FUSED_DOT")
        try:
            os.environ["B200_CG_FUSED_DOT"] = "0"
            api.cg(op.ex, op.A, b, xs, 10, 0.0)
            ms0 = min(timer.run(lambda i: api.cg(op.ex, op.A, b, xs, iters, 0.0), 1) for _ in range(2))
            it0, hist0 = api.cg(op.ex, op.A, b, xs, 5000, 1e-8, history=True)
            it0b, hist0b = api.cg(op.ex, op.A, b, xs, 5000, 1e-8, history=True)
        finally:
            if prev is None:
                os.environ.pop("B200_CG_FUSED_DOT", None)
            else:
                os.environ["B200_CG_FUSED_DOT"] = prev
        import numpy as np

        def dev(h1, h2):
            m = min(len(h1), len(h2))
            return float(np.max(np.abs(np.asarray(h1[:m]) - np.asarray(h2[:m])) / np.asarray(h2[:m]))) if m else None

        res["fused_dot_ab"] = {"ms_per_iter_separate_dot": ms0 / iters, "ms_per_iter_epilogue_dot": res["ms_per_iter"],
                               "iterations_separate_dot": int(it0), "iterations_epilogue_dot": int(holder["it"]),
                               "history_max_rel_deviation": dev(hist, hist0),
                               "history_max_rel_deviation_of_two_separate_dot_runs": dev(hist0b, hist0),
                               "how": "B200_CG_FUSED_DOT=0 vs default; histories to rtol 1e-8 compared over their common length"}
    return res


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def _bind_by_topo(idx):
    """sysfs has no NUMA node for the GPU (containers often hide it): take the CPU affinity column of
    `nvidia-smi topo -m` instead (first numeric range list on the GPU's row)."""
    import re
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
    except Exception as e:
        return f"numa: unbound (topo: {type(e).__name__})"
    for line in out.splitlines():
        tok = re.split(r"\s+", re.sub(r"\x1b\[[0-9;]*m", "", line).strip())
        if tok and tok[0] == f"GPU{idx}":
            for t in tok[1:]:
                if re.fullmatch(r"\d+(-\d+)?(,\d+(-\d+)?)*", t):
                    cpus = _parse_cpulist(t) & os.sched_getaffinity(0)
                    if not cpus:
                        return f"numa: topo affinity {t} has no allowed cpus"
                    if cpus == os.sched_getaffinity(0):
                        return f"numa: topo affinity {t} = every allowed cpu (one socket visible)"
                    os.sched_setaffinity(0, cpus)
                    return f"numa: bound by nvidia-smi topo to cpus {t} ({len(cpus)} allowed)"
    return "numa: unbound (no affinity column in nvidia-smi topo -m)"


def bind_near_gpu(local_rank):
    """N > 1: run this rank (and first-touch its pinned buffers) on the CPUs of the NUMA node its
    GPU hangs off, so host<->device copies do not cross the socket link.  Returns a note for `config`."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[local_rank]) if vis and vis.split(",")[local_rank].isdigit() else local_rank
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(idx)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:  # nvml prints an 8-digit PCI domain, sysfs a 4-digit one
            bus = bus[4:]
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read())
        if node < 0:
            return _bind_by_topo(idx)
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = set()
            for part in f.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return f"numa: node {node} has no allowed cpus"
        os.sched_setaffinity(0, cpus)
        return f"numa: bound to node {node} ({len(cpus)} cpus)"
    except Exception as e:  # no sysfs / nvml: run unbound
        return f"numa: unbound ({type(e).__name__})"


def run_b200(args):
    import torch
    from spmv_b200 import api

    rank = int(os.environ.get("RANK", "0"))
    size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert size == args.gpus or (size == 1 and args.gpus == 1), f"--gpus {args.gpus} but WORLD_SIZE={size}"
    assert torch.cuda.is_available(), "bench.py needs a GPU: there is no CPU path"
    dev = torch.device(f"cuda:{local_rank}")
    numa = bind_near_gpu(local_rank) if size > 1 else "numa: unbound (one rank)"
    torch.cuda.set_device(dev)
    if size > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        # setup collectives of the host layer (ranges, ghost counts, index lists) ride on gloo;
        # the data path (halo payload, CG scalars) is NCCL inside libb200spmv.so
        api.init_world_from_torch(dist.new_group(backend="gloo"))
    ex = api.B200Executor.create(local_rank)
    ex.use_torch_stream()
    timer = Timer(size, dev)
    peak, peak_src = peak_hbm()
    nz = args.nz
    if args.strong:
        assert NZ_PER_GPU % size == 0, "--strong needs N to divide 256"
        nz = NZ_PER_GPU // size

    # ---- headline: BASELINE configs[1], symmetric-storage 7-point SpMV -------------------
    op = Operator(api, ex, dev, rank, size, nz, 7, True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    head = measure_spmv(op, timer, args.steps, args.warmup, peak, with_e2e=True)
    head_parity = laplace_parity(op, rank, size, nz, 7, dev)
    # keep the GPUs busy ~1 s longer so the 100 ms sampler sees the loaded clocks (same count on
    # every rank: the step time is already the max over ranks)
    for _ in range(int(min(1000.0 / max(head["ms_per_step"], 1e-3), 20000))):
        op.step()
    torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    head_launches = op.launches_per_step * args.steps
    cfg = {"workload": f"BASELINE configs[1]: symmetric-storage fp64 CSR SpMV (L2GMap::update + Matrix::mult), 3-D "
                       f"7-point Laplacian {NX}x{NY}x{nz} rows per GPU, 1-D slab partition over z",
           "rows_per_gpu": op.rows, "stored_nnz_per_gpu": op.nnz_local + op.nnz_remote, "nnz_full_per_gpu": op.nnz_full(),
           "n_gpus": size, "api": "spmv_b200.api (create_laplace3d, L2GMap.update, Matrix.mult) -> libspmv_b200.so -> C ABI",
           "l2": "inputs (1.07 GB streamed per step, x and y 134 MB each) exceed the 126 MB L2; no explicit flush",
           "GFLOPs": head["GFLOPs"], "host": numa,
           "e2e_api": "Matrix.mult_host: pinned host x in, pinned host y out, copies pipelined against tile ranges"}
    op.close()
    del op

    also = {}
    if not args.no_extras:
        # ---- general 27-point SpMV (the case the roofline target is quoted on) -----------
        op27 = Operator(api, ex, dev, rank, size, nz, 27, False)
        g = measure_spmv(op27, timer, 30, 5, peak, with_e2e=True)
        also["general_27pt_spmv"] = {"parity": laplace_parity(op27, rank, size, nz, 27, dev), "GBps": g["GBps"], "ms_per_step": g["ms_per_step"], "GFLOPs": g["GFLOPs"],
                                     "kernel_ms_avg": g["kernel_ms_avg"], "kernel_ms_min": g["kernel_ms_min"],
                                     "kernel_ms_median": g["kernel_ms_median"], "kernel_GBps": g["kernel_GBps"],
                                     "frac_of_hbm_peak": g["frac_of_hbm_peak"], "e2e": g["e2e"],
                                     "algorithmic_bytes_per_gpu": op27.algorithmic_bytes()}
        # ---- CG on the 27-point operator (configs 3/4 shape) -----------------------------
        also["cg_27pt"] = measure_cg(api, op27, timer, dev, size)
        op27.close()
        del op27
        # ---- CG on the symmetric-storage 7-point operator --------------------------------
        ops = Operator(api, ex, dev, rank, size, nz, 7, True)
        also["cg_sym_7pt"] = measure_cg(api, ops, timer, dev, size)
        ops.close()
        del ops
        if size == 1 and not args.no_library_context:
            # what the reference's CUDA backend would run on this GPU: cusparseSpMV + cuBLAS BLAS-1 (library code)
            also["library_context"] = {"cusparse_general_27pt": measure_cusparse(dev, nz),
                                       "cusparse_powerlaw_3M": measure_cusparse_powerlaw(api, dev),
                                       "cublas_cg_vector_ops": measure_cublas_blas1(dev, NX * NY * nz)}
        # ---- config 1: the reference's own demo loop on create_A(1e6) --------------------------
        also["config1_createA"] = measure_config1(api, ex, dev, rank, size, timer)
        # ---- config 5: power-law rows through the PETSc reader --------------------------
        if args.config5_rows > 0:
            also["powerlaw_petsc"] = measure_config5(api, ex, dev, rank, size, timer, peak, args.config5_rows)

    if rank != 0:
        return
    line = {
        "metric": "spmv_algorithmic_GBps", "value": head["GBps"], "unit": "GB/s", "n_gpus": size, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": head["ms_per_step"], "higher_is_better": True,
        "scaling": "strong" if args.strong else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg, "e2e": head["e2e"],
        "gpu_launches": head_launches, "clocks": clocks, "parity": head_parity,
        "roofline": {"bound": "hbm", "achieved": head["kernel_GBps"], "peak": peak, "unit": "GB/s",
                     "frac": head["frac_of_hbm_peak"], "traffic": ncu_traffic("sym7")[0], "traffic_source": ncu_traffic("sym7")[1],
                     "kernel": "spmv_pipe_kernel<double, SYM, ROW4S> (local block; the event pair also spans the "
                               "zero-fill of y, zero_fill_desc_kernel, that precedes it)",
                     "kernel_ms_avg": head["kernel_ms_avg"], "kernel_ms_min": head["kernel_ms_min"],
                     "algorithmic_bytes": head["local_kernel_bytes"],
                     "peak_source": peak_src, "frac_of_8TBps_nominal": head["kernel_GBps"] / 8000.0},
        "also": also,
    }
    if size == 1 and not args.no_extras:
        line["cpu_baseline"] = cpu_baseline()
    print(json.dumps(line), file=_JSON_OUT, flush=True)


def measure_cusparse(dev, nz, steps=20, warmup=3):
    """cusparseSpMV on the general 27-point operator via torch.sparse (int32 indices, fp64), CUDA events.  Library
    code, opt-in, for context only: the reference's CUDA backend calls the same routine."""
    try:
        import torch
        from spmv_b200 import kernels as K
        from spmv_b200 import synth
        g = K.laplace3d(NX, NY, nz, 0, nz, 27, 0, dev)
        rp, ci, v = g["local"]
        n = NX * NY * nz
        A = torch.sparse_csr_tensor(rp, ci, v, size=(n, n))
        x = torch.from_numpy(synth.bump_vector(n, 0, n)).to(dev)
        for _ in range(warmup):
            y = A.matmul(x)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            y = A.matmul(x)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        nbytes = bytes_general(int(ci.numel()), 0, n, n, False)
        del A, y, g
        return {"ms_per_step": ms, "GBps": nbytes / ms / 1e6, "how": "torch.sparse_csr_tensor(int32, fp64).matmul -> cusparseSpMV"}
    except Exception as e:  # opt-in context measurement: never fail the bench over it
        return {"error": str(e)[:200]}


def measure_cusparse_powerlaw(api, dev, rows=3_000_000, steps=20):
    """cusparseSpMV (torch.sparse CSR) and this repo's kernel on the same 3 M-row power-law matrix (config 5's row
    shape): the library number the irregular-row walk is up against.  Context only."""
    try:
        import numpy as np
        import torch
        from spmv_b200 import kernels as K
        from spmv_b200 import synth
        rp, col, val = api.powerlaw_rows(rows, 0, rows)
        rp_d = torch.from_numpy(rp.astype(np.int32)).to(dev)
        ci_d = torch.from_numpy(col.astype(np.int32)).to(dev)
        v_d = torch.from_numpy(val).to(dev)
        x = torch.from_numpy(synth.bump_vector(rows, 0, rows)).to(dev)
        y = torch.empty(rows, dtype=torch.float64, device=dev)
        A = torch.sparse_csr_tensor(rp_d, ci_d, v_d, size=(rows, rows))
        plan = K.CsrPlan(rows, rows, rp_d, False, nnz=int(rp[-1]))

        def timed(fn):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / steps

        ms_lib = timed(lambda: A.matmul(x))
        ms_own = timed(lambda: K.spmv(plan, rp_d, ci_d, v_d, x, y))
        err = float((A.matmul(x) - y).norm() / y.norm())
        nbytes = bytes_general(int(rp[-1]), 0, rows, rows, False)
        plan.close()
        return {"rows": rows, "nnz": int(rp[-1]), "cusparse_ms": ms_lib, "cusparse_GBps": nbytes / ms_lib / 1e6,
                "this_repo_ms": ms_own, "this_repo_GBps": nbytes / ms_own / 1e6, "rel_diff": err}
    except Exception as e:
        return {"error": f"{type(e).__name__}: {e}"[:200]}


def measure_cublas_blas1(dev, m, iters=20):
    """The vector half of one iteration of the reference's CUDA CG (cuda/cg.cuda.cu:88-150): cublasDdot x2,
    cublasDaxpy x3, cublasDscal x1 with device pointer mode, on vectors of m doubles.  Library code, context
    only: what this repo's three vector kernels replace."""
    try:
        import ctypes as C
        import glob
        import torch
        cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "lib", "libcublas.so*")) \
            + glob.glob(os.path.join(os.path.dirname(os.path.dirname(torch.__file__)), "nvidia", "cublas", "lib", "libcublas.so*")) \
            + glob.glob("/usr/local/cuda/lib64/libcublas.so*")
        if not cands:
            return {"error": "libcublas not found"}
        L = C.CDLL(sorted(cands)[0])
        h = C.c_void_p()
        assert L.cublasCreate_v2(C.byref(h)) == 0
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        assert L.cublasSetStream_v2(h, st) == 0
        assert L.cublasSetPointerMode_v2(h, 1) == 0  # CUBLAS_POINTER_MODE_DEVICE (cg.cuda.cu:62-63)
        p, ap, x, r = (torch.rand(m, dtype=torch.float64, device=dev) for _ in range(4))
        sc = torch.tensor([1e-3, -1e-3, 0.5, 0.0, 0.0], dtype=torch.float64, device=dev)
        P = lambda t, off=0: C.c_void_p(t.data_ptr() + 8 * off)  # noqa: E731
        n, one = C.c_int(m), C.c_int(1)

        def it():
            L.cublasDdot_v2(h, n, P(p), one, P(ap), one, P(sc, 3))      # p.Ap        :109
            L.cublasDaxpy_v2(h, n, P(sc, 0), P(p), one, P(x), one)       # x += a p    :126
            L.cublasDaxpy_v2(h, n, P(sc, 1), P(ap), one, P(r), one)      # r -= a Ap   :121
            L.cublasDdot_v2(h, n, P(r), one, P(r), one, P(sc, 4))        # r.r         :131
            L.cublasDscal_v2(h, n, P(sc, 2), P(p), one)                  # p *= beta   :148
            L.cublasDaxpy_v2(h, n, P(sc, 2), P(r), one, P(p), one)       # p += r      :150

        for _ in range(3):
            it()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            it()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        L.cublasDestroy_v2(h)
        return {"ms_per_iteration": ms, "vector_passes": 14, "GBps": 14 * 8 * m / ms / 1e6,
                "how": "cublasDdot x2 + cublasDaxpy x3 + cublasDscal x1, device pointer mode, 6 library launches"}
    except Exception as e:  # context measurement: never fail the bench over it
        return {"error": f"{type(e).__name__}: {e}"[:200]}


def ncu_traffic(key):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the newest committed
    profiles/r*_traffic.json (written by profiles/tools/ncu_summary.py --traffic-json from an `ncu --set full`
    capture of this bench command); None when no capture of that kernel is on file."""
    import glob
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")), reverse=True):
        try:
            with open(f) as fh:
                d = json.load(fh)
            if key in d:
                return d[key]["dram_bytes"], os.path.basename(f)
        except Exception:
            continue
    return None, None


# ----------------------------------------------------------------------------
# CPU legs (bench_cpu_legs.py): the reference's own CPU implementation of the path on the box's host cores.
# cpu_baseline / --impl reference = its OpenMP kernel on the FULL headline operator (symmetric-storage
# 7-point 256^3, same config as the GPU arm); `legs` carries the CPU twin of every GPU leg in `also`.
# ----------------------------------------------------------------------------
CPU_WORKLOAD = (f"BASELINE configs[1]: symmetric-storage fp64 CSR SpMV, 3-D 7-point Laplacian {NX}x{NY}x{NZ_PER_GPU} "
                "(the whole 1-GPU workload); each step = one SpMV of the reference's CPU kernel")


def cpu_baseline(budget_s=8.0, with_legs=True):
    import bench_cpu_legs as CL
    h = CL.omp_spmv(7, True, nz=NZ_PER_GPU, budget_s=budget_s)
    # value = from the MEDIAN SpMV time: the box's host memory system is shared, and a burst of someone else's
    # traffic should not make the CPU look slower than it is (mean and minimum are recorded beside it)
    out = {"value": h["GBps_median"], "unit": "GB/s", "cores": h["cores"], "kind": h["kind"], "ms_per_spmv": h["ms_median"],
           "ms_mean": h["ms_per_spmv"], "ms_min": h["ms_min"], "GBps_mean": h["GBps"],
           "GFLOPs": h["GFLOPs"], "same_config": True,
           "sample": f"{h['what']}; median of {h['spmvs_timed']} SpMVs of the full operator ({h['rows']} rows)"}
    if with_legs:
        out["legs"] = CL.all_legs()
    return out


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import bench_cpu_legs as CL
    steps = max(1, min(args.steps, 200))
    warm = max(1, min(args.warmup, 20))
    h = CL.omp_spmv(7, True, nz=NZ_PER_GPU, steps=steps, warmup=warm)
    v = h["GBps"]
    print(json.dumps({
        "impl": "reference", "metric": "spmv_algorithmic_GBps", "value": v, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": h["ms_per_spmv"], "ms_median": h["ms_median"], "ms_min": h["ms_min"],
        "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": CPU_WORKLOAD, "rows": h["rows"], "stored_nnz": h["stored_nnz"], "same_config_as_gpu_arm": True},
        "cpu_baseline": {"value": v, "unit": "GB/s", "cores": h["cores"], "kind": h["kind"],
                         "sample": f"{h['what']}; {steps} SpMVs of the full operator",
                         "legs": CL.all_legs() if not args.no_extras else None},
        "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


_JSON_OUT = sys.stdout


def main():
    global _JSON_OUT
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    # stdout carries exactly one JSON line: anything a library prints to fd 1 (NCCL's version
    # banner, for one) goes to stderr instead
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    run_b200(args)
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass


if __name__ == "__main__":
    main()
