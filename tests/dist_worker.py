"""Multi-process worker for the distributed tests (launched by torch.distributed.run).

mode cpu : gloo, host executor -- the N>1 host logic (halo plan, create_matrix split,
           read_petsc, create_A) over the torch.distributed transport, checked against
           the reference's own code run with thread-ranks (oracle/_ref) or, where that
           library is absent, against the committed golden fixtures.
mode gpu : nccl, one rank per GPU -- update + mult and CG through the whole stack,
           checked against the same oracle.
Exit code 0 = all checks passed on this rank.
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from oracle import ref_api as R  # noqa: E402
from spmv_b200 import api, synth  # noqa: E402
from tests import cases  # noqa: E402
from tests.fixtures import rel_l2  # noqa: E402

TOL_Y = 1e-12


def reference_world(size, rows, N, sym, cm, want_cg=None, x_fn=None):
    """Run the reference on `size` thread-ranks; returns per-rank dicts."""
    out = {}

    def body(rank):
        inp = cases.local_inputs(rows, N, rank, size)
        A = R.RefMatrix.create_matrix(inp["rowptr"], inp["colind"], inp["values"], inp["nrows_local"],
                                      inp["ncols_local"], [], inp["col_ghosts"], sym, cm)
        plan = A.plan(0)
        d = {"plan": plan, "rowplan": A.plan(1), "local": A.block(0), "remote": A.block(1), "info": A.info()}
        n_pad = plan["local_size"] + plan["num_ghosts"]
        x = np.zeros(n_pad)
        x[: plan["local_size"]] = x_fn(inp["r0"], inp["r1"])
        y = np.zeros(inp["nrows_local"])
        A.update(x)
        A.mult(x, y)
        d["x"], d["y"] = x, y
        if want_cg:
            ones = np.zeros(n_pad)
            ones[: plan["local_size"]] = 1.0
            b = np.zeros(inp["nrows_local"])
            A.update(ones)
            A.mult(ones, b)
            d["b"] = b
            d["cg_it"], d["cg_x"] = A.cg(b, *want_cg)
        out[rank] = d
        A.close()

    R.run_ranks(size, body)
    return out


def main():
    mode = sys.argv[1]
    rank, size = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    if mode == "gpu":
        dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
        torch.cuda.set_device(dev)
        dist.init_process_group("nccl", device_id=dev)
        setup_group = dist.new_group(backend="gloo")
    else:
        dist.init_process_group("gloo")
        setup_group = None
    api.init_world_from_torch(setup_group)
    assert api.world_rank() == rank and api.world_size() == size
    ex = api.B200Executor.create(dev.index) if mode == "gpu" else api.ReferenceExecutor.create()
    have_ref = R.available()
    failures = []

    def check(cond, what):
        if not cond:
            failures.append(what)

    golden_path = os.path.join(ROOT, "tests", "golden", "plans_and_blocks.npz")
    golden = np.load(golden_path) if os.path.exists(golden_path) else None

    def check_golden(A, name, sym, cm, tag):
        """Plans and blocks against the committed output of the reference's own code."""
        key = f"{name}/s{size}/sym{int(sym)}/cm{cm}/r{rank}"
        if golden is None or f"{key}/non_zeros" not in golden.files:
            return
        plan = A.col_map().plan()
        for k in cases.PLAN_KEYS:
            check(np.array_equal(np.asarray(plan[k]), golden[f"{key}/plan/{k}"]), tag + f": golden plan {k} differs")
        for name, blk in (("local", A.local_block()), ("remote", A.remote_block())):
            have = f"{key}/{name}/dims" in golden.files
            check((blk is not None) == have, tag + f": golden {name} block presence differs")
            if blk is None or not have:
                continue
            for k in ("rowptr", "colind", "values"):
                check(np.array_equal(blk[k], golden[f"{key}/{name}/{k}"]), tag + f": golden {name} {k} differs")
            if blk["diagonal"] is not None:
                check(np.array_equal(blk["diagonal"], golden[f"{key}/{name}/diagonal"]), tag + ": golden diagonal differs")
        check(A.non_zeros() == int(golden[f"{key}/non_zeros"]), tag + ": golden non_zeros differs")

    problems = [("fixture", None), ("lap7", (4, 3, 6)), ("lap27", (6, 5, 8)), ("lap7", (7, 6, 8)), ("powerlaw", 4000), ("create_a", 1001)]
    combos = [(False, api.P2P_BLOCKING), (False, api.P2P_NONBLOCKING), (False, api.COLLECTIVE_BLOCKING),
              (False, api.COLLECTIVE_NONBLOCKING), (True, api.COLLECTIVE_BLOCKING), (True, api.P2P_NONBLOCKING)]
    for kind, n in problems:
        N, rows = cases.make_problem(kind, n)
        if N < 2 * size:
            continue  # (the 5-row fixture at 4+ ranks: ranks without rows are not a case the reference runs either)
        x_fn = lambda r0, r1: synth.bump_vector(N, r0, r1) + 0.1  # noqa: E731
        for sym, cm in combos:
            if sym and kind == "powerlaw":
                continue
            tag = f"{kind} size={size} sym={sym} cm={cm}"
            want_cg = (200, 1e-8) if (mode == "gpu" and kind in ("lap27", "lap7", "create_a")) else None
            ref = reference_world(size, rows, N, sym, cm, want_cg, x_fn)[rank] if have_ref else None
            inp = cases.local_inputs(rows, N, rank, size)
            A = api.Matrix.create_matrix(ex, inp["rowptr"], inp["colind"], inp["values"], inp["nrows_local"],
                                         inp["ncols_local"], [], inp["col_ghosts"], sym, cm)
            plan = A.col_map().plan()
            check_golden(A, cases.problem_name(kind, n), sym, cm, tag)
            if ref is not None:
                check(cases.same_plan(plan, ref["plan"]), tag + ": column map plan differs")
                check(cases.same_plan(A.row_map().plan(), ref["rowplan"]), tag + ": row map plan differs")
                check(cases.same_block(A.local_block(), ref["local"]), tag + ": local block differs")
                check(cases.same_block(A.remote_block(), ref["remote"]), tag + ": remote block differs")
                check(A.non_zeros() == ref["info"]["non_zeros"], tag + ": non_zeros differs")
            if mode == "cpu":
                # there is no CPU compute path: update must refuse, not fall back
                xh = np.zeros(plan["local_size"] + plan["num_ghosts"])
                try:
                    A.col_map().update(xh.ctypes.data)
                    check(False, tag + ": update ran on the host executor")
                except api.B200Error:
                    pass
            elif ref is not None:
                n_pad = plan["local_size"] + plan["num_ghosts"]
                x = torch.zeros(n_pad, dtype=torch.float64, device=dev)
                x[: plan["local_size"]] = torch.from_numpy(x_fn(inp["r0"], inp["r1"])).to(dev)
                y = torch.full((inp["nrows_local"],), float("nan"), dtype=torch.float64, device=dev)
                A.col_map().update(x)
                A.mult(x, y)
                ex.synchronize()
                check(np.array_equal(x.cpu().numpy(), ref["x"]), tag + ": ghost values differ after update")
                check(rel_l2(y.cpu().numpy(), ref["y"]) < TOL_Y or np.array_equal(y.cpu().numpy(), ref["y"]),
                      tag + f": y rel err {rel_l2(y.cpu().numpy(), ref['y'])}")
                # Matrix::mult_host: host vectors, copies pipelined (neighbours: edge pages, halo and
                # remote block first, local chunks add on top) -- same values as update + mult
                yh = np.full(inp["nrows_local"], np.nan)
                for _ in range(2):
                    A.mult_host(np.ascontiguousarray(x_fn(inp["r0"], inp["r1"])), yh)
                check(np.array_equal(yh, y.cpu().numpy()) if not sym else rel_l2(yh, y.cpu().numpy()) < TOL_Y,
                      tag + f": mult_host vs mult rel err {rel_l2(yh, y.cpu().numpy())}")
                if want_cg:
                    b = torch.from_numpy(ref["b"]).to(dev)
                    xs = torch.zeros_like(b)
                    it, hist = api.cg(ex, A, b, xs, want_cg[0], want_cg[1], history=True)
                    check(abs(it - ref["cg_it"]) <= 1, tag + f": cg iterations {it} vs {ref['cg_it']}")
                    xr = ref["cg_x"]
                    err = np.linalg.norm(xs.cpu().numpy() - xr) / max(np.linalg.norm(xr), 1e-300)
                    check(err < 1e-7, tag + f": cg x rel err {err}")
                    if not sym and kind in ("lap27", "lap7"):
                        # bit-identical SpMV + order-independent (Dot2) dot products on both sides, and the sum over
                        # the ranks runs in rank order on both sides (peer allreduce / fused step here, the thread-rank
                        # shim's MPI_Allreduce there): the GPU follows the reference's trajectory exactly at any rank count
                        check(it == ref["cg_it"] and np.array_equal(xs.cpu().numpy(), xr),
                              tag + f": cg not bit-identical to the reference (its {it} vs {ref['cg_it']}, err {err})")
                    if kind in ("lap27", "lap7") and cm == api.P2P_NONBLOCKING:
                        # every switch of host/cg.cpp away from its default: same trajectory
                        for envs in ({"B200_CG_FUSED_VEC": "1"}, {"B200_CG_FUSED_VEC": "1", "B200_CG_FUSED_PUSH": "1"},
                                     {"B200_CG_FUSED_VEC": "1", "B200_CG_FUSED_DOT_REMOTE": "0"},
                                     {"B200_CG_TAIL_ALLREDUCE": "0"}, {"B200_CG_FUSED_DOT_REMOTE": "0"},
                                     {"B200_CG_WAIT_IN_REMOTE": "0"}, {"B200_CG_GRAPH": "0"},
                                     {"B200_CG_FUSED_DOT": "1"}, {"B200_CG_FUSED_DOT": "0"}):
                            os.environ.update(envs)
                            try:
                                xs2 = torch.zeros_like(b)
                                it2, _ = api.cg(ex, A, b, xs2, want_cg[0], want_cg[1], history=True)
                            finally:
                                for k in envs:
                                    os.environ.pop(k, None)
                            if sym:
                                e2 = np.linalg.norm(xs2.cpu().numpy() - xr) / max(np.linalg.norm(xr), 1e-300)
                                check(abs(it2 - it) <= 1 and e2 < 1e-7, tag + f" {envs}: cg its {it2} vs {it}, err {e2}")
                            else:
                                check(it2 == it and torch.equal(xs2, xs), tag + f" {envs}: cg differs from the default path")
            A.close()

    # uneven partitions, a one-row rank, a rank that ghosts nothing (host logic only: plan + block split of the
    # product's create_matrix against the message-free restatement and the reference's own code)
    if mode == "cpu":
        from oracle import plan as OP
        for seed in range(4):
            inps = cases.random_uneven_inputs(seed, size)
            want = OP.plans([i["ncols_local"] for i in inps], [i["col_ghosts"] for i in inps])[rank]
            ref_blocks = {}
            if have_ref:
                def body(q, inps=inps, ref_blocks=ref_blocks):
                    i = inps[q]
                    B = R.RefMatrix.create_matrix(i["rowptr"], i["colind"], i["values"], i["nrows_local"], i["ncols_local"],
                                                  [], i["col_ghosts"], False, api.P2P_NONBLOCKING)
                    ref_blocks[q] = (B.plan(0), B.block(0), B.block(1))
                    B.close()
                R.run_ranks(size, body)
            inp = inps[rank]
            for cm in (api.COLLECTIVE_BLOCKING, api.P2P_NONBLOCKING):
                tag = f"random uneven seed={seed} size={size} cm={cm}"
                A = api.Matrix.create_matrix(ex, inp["rowptr"], inp["colind"], inp["values"], inp["nrows_local"],
                                             inp["ncols_local"], [], inp["col_ghosts"], False, cm)
                check(cases.same_plan(A.col_map().plan(), want), tag + ": plan differs from the restatement")
                if have_ref and cm == api.P2P_NONBLOCKING:
                    check(cases.same_plan(A.col_map().plan(), ref_blocks[rank][0]), tag + ": plan differs from the reference")
                    check(cases.same_block(A.local_block(), ref_blocks[rank][1]), tag + ": local block differs")
                    check(cases.same_block(A.remote_block(), ref_blocks[rank][2]), tag + ": remote block differs")
                A.close()

    # device-side builder (demos/CreateLaplace3D.h) against create_matrix on host-generated rows
    if mode == "gpu":
        nx, ny, nz = 6, 5, 8 * size if size not in (1, 2, 4, 8) else 8
        for st in (7, 27):
            N, rows = cases.make_problem("lap7" if st == 7 else "lap27", (nx, ny, nz))
            for sym, cm in combos:
                tag = f"create_laplace3d st={st} size={size} sym={sym} cm={cm}"
                inp = cases.local_inputs(rows, N, rank, size)
                A0 = api.Matrix.create_matrix(ex, inp["rowptr"], inp["colind"], inp["values"], inp["nrows_local"],
                                              inp["ncols_local"], [], inp["col_ghosts"], sym, cm)
                A1 = api.create_laplace3d(ex, nx, ny, nz, st, sym, cm)
                check(cases.same_plan(A1.col_map().plan(), A0.col_map().plan()), tag + ": column map differs")
                check(cases.same_plan(A1.row_map().plan(), A0.row_map().plan()), tag + ": row map differs")
                check(cases.same_block(A1.local_block(), A0.local_block()), tag + ": local block differs")
                check(cases.same_block(A1.remote_block(), A0.remote_block()), tag + ": remote block differs")
                check(A1.non_zeros() == A0.non_zeros() and A1.rows() == A0.rows() and A1.cols() == A0.cols(),
                      tag + ": sizes differ")
                A0.close()
                A1.close()

    # scalar allreduce of the CG loop: peer-memory kernel (when the ranks can map each other) and NCCL
    # give the same bits on every rank, repeatedly (the mailboxes are double-buffered by parity)
    if mode == "gpu":
        A = api.create_laplace3d(ex, 6, 5, 8 * size, 7, False, api.P2P_NONBLOCKING)
        cmap = A.col_map()
        for use_nccl in (False, True):
            for rep in range(5):
                v = torch.tensor([rank + 1.0, 0.25 * (rank + rep), -1e-3 * rank], dtype=torch.float64, device=dev)
                cmap.allreduce_sum(v, nccl=use_nccl)
                ex.synchronize()
                want = [sum(r + 1.0 for r in range(size)), sum(0.25 * (r + rep) for r in range(size)),
                        sum(-1e-3 * r for r in range(size))]
                got = v.cpu().numpy()
                check(np.allclose(got, want, rtol=1e-15, atol=0) and got[0] == want[0],
                      f"allreduce_sum nccl={use_nccl} rep={rep} size={size}: {got} vs {want}")
                gathered = [torch.zeros_like(v) for _ in range(size)]
                dist.all_gather(gathered, v)
                check(all(torch.equal(g, gathered[0]) for g in gathered), f"allreduce_sum nccl={use_nccl}: ranks disagree")
        A.close()

    # L2GMap::reverse_update (L2GMap.cpp:908-923): ghost values travel back to their owners and are added
    # onto vec[indexbuf[k]].  Expected values from the message-free restatement (oracle/plan.py, itself
    # pinned to the reference's reverse_update_collective in tests/test_oracle.py), three calls in a row (the
    # peer staging slots alternate by epoch), peer-memory exchange and the NCCL fallback, fp64 and fp32.
    if mode == "gpu":
        from oracle import plan as OP
        for kind, n in (("fixture", None), ("lap7", (7, 6, 8)), ("powerlaw", 4000)):
            N, rows = cases.make_problem(kind, n)
            if N < 2 * size:
                continue
            inps = [cases.local_inputs(rows, N, r, size) for r in range(size)]
            plans = OP.plans([i["ncols_local"] for i in inps], [i["col_ghosts"] for i in inps])
            for halo in ("peer", "kernel", "nccl"):  # copy-engine exchange (default), kernel-driven exchange, NCCL
                if halo != "peer":
                    os.environ["B200_HALO"] = halo
                inp = inps[rank]
                A = api.Matrix.create_matrix(ex, inp["rowptr"], inp["colind"], inp["values"], inp["nrows_local"],
                                             inp["ncols_local"], [], inp["col_ghosts"], False, api.COLLECTIVE_BLOCKING)
                os.environ.pop("B200_HALO", None)
                cmap = A.col_map()
                check(cases.same_plan(cmap.plan(), plans[rank]), f"reverse_update {kind}: plan differs from the restatement")
                for dt, tol in ((torch.float64, 1e-14), (torch.float32, 1e-6)):
                    npdt = np.float64 if dt == torch.float64 else np.float32
                    vecs = [np.random.default_rng(77 + r).standard_normal(p["local_size"] + p["num_ghosts"]).astype(npdt)
                            for r, p in enumerate(plans)]
                    x = torch.from_numpy(vecs[rank].copy()).to(dev)
                    for rep in range(3):
                        cmap.reverse_update(x)
                        OP.reverse_update(plans, vecs)
                        for v, p in zip(vecs, plans):  # the tail is not changed by the call; refresh it so every
                            v[p["local_size"]:] += 1     # epoch moves different numbers
                        x[plans[rank]["local_size"]:] += 1
                    ex.synchronize()
                    err = rel_l2(x.cpu().numpy().astype(np.float64), vecs[rank].astype(np.float64))
                    check(err < tol, f"reverse_update {kind} halo={halo} {dt}: rel err {err}")
                    # forward update of the same dtype: ghosts are copies, bit for bit
                    vecs2 = [np.random.default_rng(99 + r).standard_normal(p["local_size"] + p["num_ghosts"]).astype(npdt)
                             for r, p in enumerate(plans)]
                    x2 = torch.from_numpy(vecs2[rank].copy()).to(dev)
                    cmap.update(x2)
                    cmap.update_finalise(x2)
                    OP.update(plans, vecs2)
                    ex.synchronize()
                    check(np.array_equal(x2.cpu().numpy(), vecs2[rank]), f"update {kind} halo={halo} {dt}: ghosts differ")
                A.close()

    # fp32 through the API (Matrix<float>::create_matrix -> L2GMap::update<float> -> Matrix<float>::mult)
    if mode == "gpu":
        from oracle import kernels as OKf
        for kind, n, sym, cm in (("fixture", None, False, api.P2P_BLOCKING), ("fixture", None, True, api.COLLECTIVE_BLOCKING),
                                 ("lap7", (7, 6, 8), True, api.P2P_NONBLOCKING), ("lap27", (6, 5, 8), False, api.P2P_NONBLOCKING),
                                 ("lap27", (6, 5, 8), False, api.COLLECTIVE_BLOCKING)):
            N, rows = cases.make_problem(kind, n)
            if N < 2 * size:
                continue
            inp = cases.local_inputs(rows, N, rank, size)
            A = api.MatrixF32.create_matrix(ex, inp["rowptr"], inp["colind"], inp["values"].astype(np.float32),
                                            inp["nrows_local"], inp["ncols_local"], [], inp["col_ghosts"], sym, cm)
            cmap = A.col_map()
            xg = (synth.bump_vector(N, 0, N) + 0.1).astype(np.float32)
            x = torch.zeros(cmap.local_size() + cmap.num_ghosts(), dtype=torch.float32, device=dev)
            x[: inp["nrows_local"]] = torch.from_numpy(xg[inp["r0"]:inp["r1"]]).to(dev)
            y = torch.full((inp["nrows_local"],), float("nan"), dtype=torch.float32, device=dev)
            cmap.update(x)
            A.mult(x, y)
            ex.synchronize()
            rp, col, val = rows(0, N)  # the fp32 oracle on the global operator, this rank's rows
            want = OKf.spmv(rp.astype(np.int32), col.astype(np.int32), val.astype(np.float32), xg)[inp["r0"]:inp["r1"]]
            err = rel_l2(y.cpu().numpy().astype(np.float64), want.astype(np.float64))
            check(err < 2e-6, f"fp32 {kind} sym={sym} cm={cm} size={size}: rel err {err}")
            A.close()

    # mult_host on slabs big enough for several chunks and distinct edge pages
    if mode == "gpu":
        for st, sym, cm in ((7, True, api.P2P_NONBLOCKING), (27, False, api.P2P_NONBLOCKING),
                            (27, False, api.COLLECTIVE_NONBLOCKING), (7, False, api.P2P_BLOCKING)):
            A = api.create_laplace3d(ex, 64, 48, 40 * size, st, sym, cm)
            m, cmap = A.rows(), A.col_map()
            xh = np.random.default_rng(11 + rank).standard_normal(m)
            x = torch.zeros(cmap.local_size() + cmap.num_ghosts(), dtype=torch.float64, device=dev)
            x[:m] = torch.from_numpy(xh).to(dev)
            y = torch.full((m,), float("nan"), dtype=torch.float64, device=dev)
            cmap.update(x)
            A.mult(x, y)
            ex.synchronize()
            yh = np.full(m, np.nan)
            for _ in range(3):
                A.mult_host(xh, yh)
            err = rel_l2(yh, y.cpu().numpy())
            check(np.array_equal(yh, y.cpu().numpy()) if not sym else err < TOL_Y,
                  f"mult_host slab st={st} sym={sym} cm={cm} size={size}: rel err {err}")
            A.close()

    # read_petsc + create_A builders over the multi-process transport
    N, rows = cases.make_problem("powerlaw", 3000)
    rp, col, val = rows(0, N)
    path = os.path.join(tempfile.gettempdir(), f"spmv_b200_dist_{os.environ.get('MASTER_PORT', '0')}.petsc")
    if rank == 0:
        api.write_petsc_binary_matrix(path, N, N, rp, col, val)
    dist.barrier()
    for sym, cm in [(False, api.COLLECTIVE_BLOCKING), (False, api.P2P_NONBLOCKING)]:
        A = api.read_petsc_binary_matrix(path, ex, sym, cm)
        if have_ref:
            got = {}

            def body(r):
                M = R.RefMatrix.read_petsc(path, sym, cm)
                if r == rank:
                    got["plan"], got["local"], got["remote"] = M.plan(0), M.block(0), M.block(1)
                M.close()

            R.run_ranks(size, body)
            check(cases.same_plan(A.col_map().plan(), got["plan"]), f"read_petsc cm={cm}: plan differs")
            check(cases.same_block(A.local_block(), got["local"]), f"read_petsc cm={cm}: local block differs")
            check(cases.same_block(A.remote_block(), got["remote"]), f"read_petsc cm={cm}: remote block differs")
        A.close()
    A = api.create_A(1000, ex)
    if have_ref:
        got = {}

        def body2(r):
            M = R.RefMatrix.create_A(1000)
            if r == rank:
                got["plan"], got["local"] = M.plan(0), M.block(0)
            M.close()

        R.run_ranks(size, body2)
        check(cases.same_plan(A.col_map().plan(), got["plan"]), "create_A: plan differs")
        check(cases.same_block(A.local_block(), got["local"]), "create_A: block differs")
    A.close()
    # BASELINE configs[0] as written: demos/spmv.cpp:73-108 on create_A(1 000 000) -- warm-up, 100 x {update; mult;
    # x <- y}, final ||x||_2 -- against the reference's own loop at the same number of thread-ranks: halo lists and
    # block bit-exact, the first y and the x after 100 products bit-identical (3-entry rows, one thread per row in
    # the reference's order; every rank sums its rows in the same local-column order as the reference's rank does)
    if mode == "gpu":
        n1 = 1_000_000
        A = api.create_A(n1, ex)
        cmap = A.col_map()
        m, r0 = A.rows(), cmap.global_offset()
        x = torch.zeros(m + cmap.num_ghosts(), dtype=torch.float64, device=dev)
        x[:m] = torch.from_numpy(synth.bump_vector(n1, r0, r0 + m)).to(dev)
        y = torch.empty(m, dtype=torch.float64, device=dev)
        cmap.update(x)
        A.mult(x, y)
        y_first = y.cpu().numpy()
        for _ in range(100):
            cmap.update(x)
            A.mult(x, y)
            x[:m].copy_(y)
        ex.synchronize()
        sq = (x[:m] * x[:m]).sum().reshape(1)
        dist.all_reduce(sq)
        if have_ref:
            got = {}

            def body3(r):
                M = R.RefMatrix.create_A(n1)
                pl = M.plan(0)
                mm, rr0 = pl["local_size"], pl["global_offset"]
                xr = np.zeros(mm + pl["num_ghosts"])
                xr[:mm] = synth.bump_vector(n1, rr0, rr0 + mm)
                yr = np.zeros(mm)
                M.update(xr)
                M.mult(xr, yr)
                y0 = yr.copy()
                _, nrm = M.demo_spmv_loop(xr, yr, 100)
                if r == rank:
                    got.update(plan=pl, local=M.block(0), y0=y0, x=xr[:mm].copy(), norm=nrm)
                M.close()

            R.run_ranks(size, body3)
            check(cases.same_plan(cmap.plan(), got["plan"]), "config 1: halo plan differs")
            check(cases.same_block(A.local_block(), got["local"]), "config 1: block differs")
            check(np.array_equal(y_first, got["y0"]), "config 1: first y not bit-identical")
            check(np.array_equal(x[:m].cpu().numpy(), got["x"]), "config 1: x after 100 products not bit-identical")
            nrm = float(sq.sqrt().item())
            check(abs(nrm - got["norm"]) <= 1e-12 * got["norm"], f"config 1: final norm {nrm} vs {got['norm']}")
        A.close()
    dist.barrier()
    if rank == 0 and os.path.exists(path):
        os.remove(path)

    for f in failures[:20]:
        print(f"[rank {rank}] FAIL {f}", flush=True)
    ok = torch.tensor([0 if failures else 1])
    if mode == "gpu":
        ok = ok.to(dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("DIST_OK" if int(ok.item()) == 1 else "DIST_FAILED", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(ok.item()) == 1 else 1)


if __name__ == "__main__":
    main()
