"""CPU legs of bench.py: the REFERENCE's own CPU implementation of the path, timed on the box's host cores.

Imported only by bench.py, for its `cpu_baseline` object and its `--impl reference` arm -- the one place
outside tests/ that may execute oracle/ (oracle/_ref = the reference's sources compiled unmodified where
they lie; liboracle.so = the plain-C restatement, used only when oracle/_ref is absent).  Nothing here is
on the product path.

Legs (all on the workloads bench.py's GPU legs run, sizes stated in each record):
  sym7_omp      reference OpenMP kernel (openmp/csr_kernels.openmp.cpp:173-244), symmetric-storage 7-point
                256^3, all host threads                         <-> headline (BASELINE configs[1])
  general27_omp same kernel, general storage, 27-point 256^3    <-> also.general_27pt_spmv
  mpi_sym7      "pure MPI" analogue: R = cores thread-ranks, each with the reference's sequential kernel
                (csr_kernels.cpp:20-52) behind Matrix::mult, plus L2GMap::update, in the timed loop of
                demos/spmv.cpp:84-103                           <-> headline, the reference's MPI path
  cg27_mpi      the reference's cg.cpp:21-98 at R thread-ranks on a 256x256xZ slab of the 27-point operator
                                                               <-> also.cg_27pt
  cg_sym7_mpi   same on the symmetric-storage 7-point operator  <-> also.cg_sym_7pt
  config1_mpi   BASELINE configs[0] itself: demos/spmv.cpp loop on create_A(1e6) at 2 thread-ranks   <-> also.config1
  powerlaw_omp  reference OpenMP kernel, general storage, power-law rows (3 M-row twin of config 5)
                                                               <-> also.powerlaw_petsc
"""
import os
import time

import numpy as np

NX = NY = 256


def cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def bytes_general(nnz, rows, ncols):
    return 12 * nnz + 4 * (rows + 1) + 8 * ncols + 8 * rows


def bytes_symmetric(nnz_lower, rows):
    return 12 * nnz_lower + 4 * (rows + 1) + 24 * rows


def _time_loop(run, budget_s, max_n, min_n=3, each=None):
    run()
    t0, n = time.perf_counter(), 0
    while True:
        t1 = time.perf_counter()
        run()
        n += 1
        dt = time.perf_counter() - t0
        if each is not None:
            each.append(time.perf_counter() - t1)
        if n >= min_n and (dt > budget_s or n >= max_n):
            return dt / n, n


def omp_spmv(stencil, symmetric, nz=256, budget_s=6.0, steps=None, warmup=1):
    """The reference's OpenMP CSRSpMV<double>::run on the 256x256xnz Laplacian, OMP_NUM_THREADS = all cores."""
    from oracle import kernels as OK
    from spmv_b200 import synth
    rows = NX * NY * nz
    g = OK.laplace3d(NX, NY, nz, 0, nz, stencil, symmetric)
    rp, ci, v = g[0], g[1], g[2]
    diag = g[3] if symmetric else None
    nthreads = cores()
    x = synth.bump_vector(rows, 0, rows)
    y = np.zeros(rows)
    if OK.have_ref():
        op = OK.RefOmp(rows, rows, rp, ci, v, diagonal=diag, symmetric=symmetric, threads=nthreads)
        kind, run = "reference", (lambda: op.run(x, y))
    else:  # the plain-C port, one thread
        kind, nthreads = "port", 1
        if symmetric:
            run = lambda: OK.symv(rp, ci, v, diag, x)  # noqa: E731
        else:
            run = lambda: OK.spmv(rp, ci, v, x)  # noqa: E731
    each = []  # per-SpMV wall times: the host's memory system is shared with other tenants of the box, so the
    #            median and the minimum are recorded beside the mean (the mean is what "time K steps" means)
    if steps is None:
        sec, n = _time_loop(run, budget_s, 200, each=each)
    else:
        for _ in range(max(1, warmup)):
            run()
        t0 = time.perf_counter()
        for _ in range(steps):
            t1 = time.perf_counter()
            run()
            each.append(time.perf_counter() - t1)
        sec, n = (time.perf_counter() - t0) / steps, steps
    nbytes = bytes_symmetric(len(v), rows) if symmetric else bytes_general(len(v), rows, rows)
    nnz_full = 2 * len(v) + rows if symmetric else len(v)
    return {"GBps": nbytes / sec / 1e9, "ms_per_spmv": sec * 1e3, "GFLOPs": 2 * nnz_full / sec / 1e9, "cores": nthreads,
            "ms_median": float(np.median(each)) * 1e3, "ms_min": float(np.min(each)) * 1e3,
            "GBps_median": nbytes / float(np.median(each)) / 1e9,
            "kind": kind, "rows": rows, "stored_nnz": int(len(v)), "spmvs_timed": n,
            "what": (f"{'symmetric-storage' if symmetric else 'general'} {stencil}-point Laplacian {NX}x{NY}x{nz}, "
                     + ("reference openmp/csr_kernels.openmp.cpp, OMP_NUM_THREADS=%d" % nthreads if kind == "reference"
                        else "oracle C port, 1 thread"))}


def _rank_inputs(nzg, stencil, rank, size):
    """One thread-rank's create_matrix arguments for the 256x256xnzg Laplacian (whole planes per rank)."""
    from oracle import kernels as OK
    from spmv_b200 import synth
    plane = NX * NY
    rng = synth.owner_ranges(size, plane * nzg)
    r0, r1 = int(rng[rank]), int(rng[rank + 1])
    assert r0 % plane == 0 and r1 % plane == 0, "ranks own whole planes"
    rp, gcol, val = OK.laplace3d(NX, NY, nzg, r0 // plane, r1 // plane, stencil)
    lcol, ghosts = synth.localize_columns(gcol.astype(np.int64), r0, r1)
    return rp, lcol, val, r0, r1, ghosts


def _ranks_for(nzg, want):
    """Largest rank count <= want that gives every rank whole planes (owner_ranges of nzg planes)."""
    r = max(1, min(want, nzg))
    while nzg % r:
        r -= 1
    return r


def mpi_spmv(stencil=7, symmetric=True, nz=256, n_apply=20):
    """demos/spmv.cpp:73-108 with the reference's own L2GMap / Matrix / sequential kernel at R thread-ranks."""
    from oracle import ref_api as R
    from spmv_b200 import synth
    if not R.available():
        return {"unavailable": "oracle/_ref/libspmv_ref.so not built"}
    size = _ranks_for(nz, cores())
    n_glob = NX * NY * nz
    out = {}

    def body(rank):
        rp, lcol, val, r0, r1, ghosts = _rank_inputs(nz, stencil, rank, size)
        A = R.RefMatrix.create_matrix(rp, lcol, val, r1 - r0, r1 - r0, [], ghosts, symmetric, 2)  # collective_blocking
        pl, info = A.plan(0), A.info()
        x = np.zeros(pl["local_size"] + pl["num_ghosts"])
        x[: r1 - r0] = synth.bump_vector(n_glob, r0, r1)
        y = np.zeros(r1 - r0)
        t, nrm = A.demo_spmv_loop(x, y, n_apply)
        out[rank] = (t, nrm, info["local_nnz"], info["remote_nnz"], pl["num_ghosts"])
        A.close()

    t0 = time.perf_counter()
    R.run_ranks(size, body)
    wall = time.perf_counter() - t0
    upd = max(o[0]["update"] for o in out.values()) / n_apply
    mul = max(o[0]["mult"] for o in out.values()) / n_apply
    if symmetric:
        nbytes = sum(bytes_symmetric(o[2], (NX * NY * nz) // size) + 12 * o[3] + 8 * o[4] for o in out.values())
    else:
        nbytes = sum(bytes_general(o[2] + o[3], (NX * NY * nz) // size, (NX * NY * nz) // size + o[4]) for o in out.values())
    return {"ranks": size, "cores": cores(), "kind": "reference", "ms_update_max": upd * 1e3, "ms_mult_max": mul * 1e3,
            "ms_per_step": (upd + mul) * 1e3, "GBps": nbytes / (upd + mul) / 1e9, "n_apply": n_apply,
            "final_x_norm": out[0][1], "setup_and_run_s": wall,
            "what": f"demos/spmv.cpp loop (update; mult; x<-y, min/max-over-ranks timing of :120-130), {stencil}-point "
                    f"{NX}x{NY}x{nz}, {'symmetric' if symmetric else 'general'} storage, {size} thread-ranks with the "
                    f"reference's L2GMap::update + sequential csr_kernels.cpp, collective_blocking"}


def mpi_cg(stencil, symmetric, nz=64, iters=20):
    """The reference's spmv::cg (cg.cpp:21-98) at R thread-ranks, b = A*1, fixed iteration count (rtol 0)."""
    from oracle import ref_api as R
    if not R.available():
        return {"unavailable": "oracle/_ref/libspmv_ref.so not built"}
    size = _ranks_for(nz, cores())
    out = {}

    def body(rank):
        rp, lcol, val, r0, r1, ghosts = _rank_inputs(nz, stencil, rank, size)
        A = R.RefMatrix.create_matrix(rp, lcol, val, r1 - r0, r1 - r0, [], ghosts, symmetric, 2)
        pl = A.plan(0)
        ones = np.zeros(pl["local_size"] + pl["num_ghosts"])
        ones[: r1 - r0] = 1.0
        b = np.zeros(r1 - r0)
        A.update(ones)
        A.mult(ones, b)
        A.cg(b, 2, 0.0)  # warm-up
        t0 = time.perf_counter()
        it, _ = A.cg(b, iters, 0.0)
        out[rank] = (time.perf_counter() - t0, it)
        A.close()

    R.run_ranks(size, body)
    sec = max(o[0] for o in out.values())
    its = out[0][1]
    return {"ranks": size, "cores": cores(), "kind": "reference", "iters_per_s": its / sec, "ms_per_iter": sec / its * 1e3,
            "iterations": its, "rows": NX * NY * nz,
            "what": f"reference cg.cpp at {size} thread-ranks (sequential kernel per rank, MPI shim allreduce), {stencil}-point "
                    f"{NX}x{NY}x{nz} slab ({nz}/256 of the GPU leg's rows per GPU), "
                    f"{'symmetric' if symmetric else 'general'} storage"}


def config1_mpi(n=1_000_000, ranks=2, n_apply=100):
    """BASELINE configs[0] as written: demos/spmv.cpp:73-108 on create_A(N) (demos/CreateA.cpp:31-103), general
    CSR, pure MPI at 2 ranks (thread-ranks here), ReferenceExecutor, collective_blocking."""
    from oracle import ref_api as R
    from spmv_b200 import synth
    if not R.available():
        return {"unavailable": "oracle/_ref/libspmv_ref.so not built"}
    out = {}

    def body(rank):
        A = R.RefMatrix.create_A(n)
        pl = A.plan(0)
        r0, m = pl["global_offset"], pl["local_size"]
        x = np.zeros(m + pl["num_ghosts"])
        x[:m] = synth.bump_vector(n, r0, r0 + m)
        y = np.zeros(m)
        t, nrm = A.demo_spmv_loop(x, y, n_apply)
        out[rank] = (t, nrm, A.info()["local_nnz"])
        A.close()

    R.run_ranks(ranks, body)
    t = {k: max(o[0][k] for o in out.values()) / n_apply for k in ("update", "mult", "copy")}
    nnz = sum(o[2] for o in out.values())
    step = t["update"] + t["mult"] + t["copy"]
    return {"ranks": ranks, "cores": cores(), "kind": "reference", "us_update_max": t["update"] * 1e6,
            "us_mult_max": t["mult"] * 1e6, "us_copy_max": t["copy"] * 1e6, "us_per_step": step * 1e6,
            "GBps": bytes_general(nnz, n, n) / (t["update"] + t["mult"]) / 1e9, "final_x_norm": out[0][1], "n_apply": n_apply,
            "rows": n, "what": f"demos/spmv.cpp loop on create_A({n}), {ranks} thread-ranks, reference L2GMap::update "
                               "(collective_blocking) + sequential csr_kernels.cpp + copy, max over ranks per phase"}


def powerlaw_omp(rows=3_000_000, budget_s=4.0):
    from oracle import kernels as OK
    from spmv_b200 import api, synth
    rp, col, val = api.powerlaw_rows(rows, 0, rows)
    rp, ci = rp.astype(np.int32), col.astype(np.int32)
    x = synth.bump_vector(rows, 0, rows)
    y = np.zeros(rows)
    nthreads = cores()
    if OK.have_ref():
        op = OK.RefOmp(rows, rows, rp, ci, val, threads=nthreads)
        kind, run = "reference", (lambda: op.run(x, y))
    else:
        kind, nthreads, run = "port", 1, (lambda: OK.spmv(rp, ci, val, x))
    sec, n = _time_loop(run, budget_s, 100)
    return {"GBps": bytes_general(len(val), rows, rows) / sec / 1e9, "ms_per_spmv": sec * 1e3, "cores": nthreads,
            "kind": kind, "rows": rows, "nnz": int(len(val)), "spmvs_timed": n,
            "what": f"power-law rows, {rows}-row twin of config 5, general storage, reference OpenMP kernel"}


def all_legs(light=False):
    """Every CPU twin of the GPU legs; each leg is bounded to a few seconds of timed work."""
    legs = {}
    for name, fn in (("config1_mpi", lambda: config1_mpi(100_000 if light else 1_000_000)),
                     ("general27_omp", lambda: omp_spmv(27, False, nz=64 if light else 256, budget_s=4.0)),
                     ("mpi_sym7", lambda: mpi_spmv(7, True, nz=64 if light else 256, n_apply=10 if light else 20)),
                     ("cg27_mpi", lambda: mpi_cg(27, False, nz=32 if light else 64, iters=10 if light else 20)),
                     ("cg_sym7_mpi", lambda: mpi_cg(7, True, nz=32 if light else 64, iters=10 if light else 20)),
                     ("powerlaw_omp", lambda: powerlaw_omp(300_000 if light else 3_000_000))):
        t0 = time.perf_counter()
        try:
            legs[name] = fn()
        except Exception as e:  # a CPU context leg must not take the GPU numbers down with it
            legs[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
        legs[name]["leg_wall_s"] = time.perf_counter() - t0
    return legs
