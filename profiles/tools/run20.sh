#!/bin/bash
# r02, 2 GPUs: two-channel halo validation
mkdir -p gpurun_out
true
true
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --steps 30 --warmup 3 $2 > gpurun_out/$3.json 2> gpurun_out/$3.err; echo "$3 rc=$?"; tail -2 gpurun_out/$3.err | cut -c1-300; }
run 29511 "" r02_bench_n2
run 29512 "--strong --config5-rows 0" r02_bench_n2_strong
python - <<'PY'
import json
for f in ("r02_bench_n2","r02_bench_n2_strong"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "no json", e); continue
    a=d["also"]
    print(f, round(d["value"]), d["ms_per_step"], d["parity"]["rel_l2"], "e2e", d["e2e"]["ms_per_step"], d["e2e"].get("pcie_floor_ms"))
    print(" 27pt", a["general_27pt_spmv"]["ms_per_step"], a["general_27pt_spmv"]["parity"]["rel_l2"])
    for k in ("cg_27pt","cg_sym_7pt"):
        print(" ",k, a[k]["ms_per_iter_runs"], a[k]["halo_mode"], {x:round(y,1) for x,y in a[k]["phases"].items() if x!="how"}, a[k]["solve_rtol_1e-8"]["iterations"], round(a[k]["solve_rtol_1e-8"]["iters_per_s"]))
    if "powerlaw_petsc" in a: print(" pl", a["powerlaw_petsc"]["ms_per_step"], a["powerlaw_petsc"]["parity"]["max_rel_err"])
    print(" c1", a["config1_createA"]["us_per_step"], a["config1_createA"]["halo_mode"])
PY
