#!/bin/bash
mkdir -p gpurun_out
SECONDS=0
timeout 1500 python bench.py > gpurun_out/r02_bench_n1_b.json 2> gpurun_out/r02_bench_n1_b.err; echo "bench rc=$? wall ${SECONDS}s"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_bench_n1_b.json"))
a=d["also"]
print(round(d["value"]), d["ms_per_step"], d["roofline"]["frac"], d["parity"]["rel_l2"], "e2e", d["e2e"])
print("27pt", a["general_27pt_spmv"]["ms_per_step"], a["general_27pt_spmv"]["frac_of_hbm_peak"])
for k in ("cg_27pt","cg_sym_7pt"):
    print(k, a[k]["ms_per_iter_runs"], {x:round(y,1) for x,y in a[k]["phases"].items() if x!="how"}, a[k]["solve_rtol_1e-8"])
print("pl", a["powerlaw_petsc"]["ms_per_step"], a["powerlaw_petsc"]["local_kernel_frac_of_hbm_peak"])
print("c1", a["config1_createA"])
print("lib", a["library_context"])
c=d["cpu_baseline"]; print("cpu", c["value"], c["ms_per_spmv"], c["cores"])
for k,v in c["legs"].items(): print(" ",k,{x:y for x,y in v.items() if x!="what"})
PY
CMD="python bench.py --steps 3 --warmup 3 --no-extras"
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02_launches_bench_sym7.csv $CMD > gpurun_out/r02_ncu_ll.log 2>&1
echo "ncu launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:spmv_pipe -s 6 -c 2 -f -o gpurun_out/r02_sym7 $CMD > gpurun_out/r02_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/r02_ncu_full.log
