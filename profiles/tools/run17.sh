#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu_b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu_b.log
tail -6 gpurun_out/r02_pytest_gpu_b.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
/usr/bin/time -v timeout 1200 python bench.py > gpurun_out/r02_bench_n1_b.json 2> gpurun_out/r02_bench_n1_b.err; echo "bench rc=$?"
grep -E "Elapsed|Maximum resident" gpurun_out/r02_bench_n1_b.err
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
timeout 600 python bench.py --impl reference --steps 20 --warmup 2 --no-extras | cut -c1-600
