#!/bin/bash
# r02 final 1-GPU validation: the new test alone in a fresh process, the full GPU suite, smoke, bench (both arms)
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02_pytest_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r02_pytest_noalloc_alone.log
tail -3 gpurun_out/r02_pytest_noalloc_alone.log | cut -c1-300
python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu_c.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu_c.log
tail -6 gpurun_out/r02_pytest_gpu_c.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
S0=$SECONDS
timeout 1200 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc=$? in $((SECONDS-S0)) s"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_bench_n1.json"))
a=d["also"]
print(round(d["value"]), d["ms_per_step"], d["roofline"], d["parity"]["rel_l2"], "e2e", d["e2e"])
print("27pt", a["general_27pt_spmv"]["ms_per_step"], a["general_27pt_spmv"]["frac_of_hbm_peak"])
for k in ("cg_27pt","cg_sym_7pt"):
    print(k, a[k]["ms_per_iter_runs"], {x:round(y,1) for x,y in a[k]["phases"].items() if x!="how"}, a[k]["solve_rtol_1e-8"])
print("pl", a["powerlaw_petsc"]["ms_per_step"], a["powerlaw_petsc"]["local_kernel_frac_of_hbm_peak"])
print("c1", a["config1_createA"])
print("lib", a["library_context"])
c=d["cpu_baseline"]; print("cpu", c["value"], c.get("ms_per_spmv"), c["cores"])
print("clocks", d["clocks"], "launches", d["gpu_launches"])
PY
S0=$SECONDS
timeout 600 python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/r02_bench_ref_n1.json 2> gpurun_out/r02_bench_ref_n1.err; echo "ref rc=$? in $((SECONDS-S0)) s"
cut -c1-500 gpurun_out/r02_bench_ref_n1.json
