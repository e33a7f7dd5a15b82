#!/bin/bash
# r02 final, weak scaling (bench.py default) at 8 and 4 GPUs
mkdir -p gpurun_out
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $3 "${@:4}"; }
run() { tr 600 $1 $2 bench.py --gpus $1 --steps 30 --warmup 3 $3 > gpurun_out/$4.json 2> gpurun_out/$4.err; echo "$4 rc=$?"; tail -2 gpurun_out/$4.err | cut -c1-300; }
S0=$SECONDS
run 8 29701 "" r02_bench_n8
echo "t=$((SECONDS-S0))"
run 4 29704 "--config5-rows 0" r02_bench_n4
echo "t=$((SECONDS-S0))"
python - <<'PY'
import json
for f in ("r02_bench_n8","r02_bench_n4"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "no json", e); continue
    a=d["also"]
    print(f, round(d["value"]), d["ms_per_step"], d["parity"]["rel_l2"], "e2e", d["e2e"]["ms_per_step"], d["e2e"].get("pcie_floor_ms"))
    print(" 27pt", a["general_27pt_spmv"]["ms_per_step"], a["general_27pt_spmv"]["parity"]["rel_l2"])
    for k in ("cg_27pt","cg_sym_7pt"):
        print(" ",k, a[k]["ms_per_iter_runs"], {x:round(y,1) for x,y in a[k]["phases"].items() if x!="how"}, a[k]["solve_rtol_1e-8"]["iterations"], round(a[k]["solve_rtol_1e-8"]["iters_per_s"]), a[k]["solve_rtol_1e-8"]["residual_true"])
    if "powerlaw_petsc" in a: print(" pl", a["powerlaw_petsc"]["ms_per_step"], a["powerlaw_petsc"]["local_kernel_frac_of_hbm_peak"], a["powerlaw_petsc"]["parity"]["max_rel_err"])
    print(" c1", a["config1_createA"]["us_per_step"])
PY
