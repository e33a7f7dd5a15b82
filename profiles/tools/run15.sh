#!/bin/bash
# r02, 2 GPUs: distributed GPU tests (kernel halo, reverse_update, fp32, fused step over peers, graph) + bench weak/strong
mkdir -p gpurun_out
export B200_CG_DEBUG=1
timeout 900 python -m pytest tests/test_dist.py -x -q -m gpu > gpurun_out/r02_pytest_dist_n2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_dist_n2.log
tail -25 gpurun_out/r02_pytest_dist_n2.log | cut -c1-300
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --steps 20 --warmup 3 $2 > gpurun_out/$3.json 2> gpurun_out/$3.err; echo "$3 rc=$?"; tail -3 gpurun_out/$3.err | cut -c1-300; }
run 29511 "" r02_bench_n2_a
run 29512 "--strong --config5-rows 0" r02_bench_n2_strong_a
python - <<'PY'
import json
for f in ("r02_bench_n2_a","r02_bench_n2_strong_a"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "no json", e); continue
    a=d["also"]
    print(f, d["value"], d["ms_per_step"], d["parity"]["rel_l2"], d["e2e"]["ms_per_step"])
    print(" 27pt", a["general_27pt_spmv"]["ms_per_step"], a["general_27pt_spmv"]["parity"]["rel_l2"])
    for k in ("cg_27pt","cg_sym_7pt"):
        print(" ",k, a[k]["ms_per_iter_runs"], a[k]["halo_mode"], a[k]["phases"], a[k]["solve_rtol_1e-8"])
    if "powerlaw_petsc" in a: print(" pl", a["powerlaw_petsc"]["ms_per_step"], a["powerlaw_petsc"]["parity"])
PY
