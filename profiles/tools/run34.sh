#!/bin/bash
# r02 final 1-GPU validation on the final tree
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02_pytest_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r02_pytest_noalloc_alone.log
tail -2 gpurun_out/r02_pytest_noalloc_alone.log | cut -c1-300
python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu.log
tail -4 gpurun_out/r02_pytest_gpu.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
S0=$SECONDS
timeout 1200 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc=$? in $((SECONDS-S0)) s"
S0=$SECONDS
timeout 600 python bench.py --strong --config5-rows 0 --no-library-context > gpurun_out/r02_bench_n1_strong.json 2> gpurun_out/r02_bench_n1_strong.err; echo "bench strong rc=$? in $((SECONDS-S0)) s"
python - <<'PY'
import json
for f in ("r02_bench_n1","r02_bench_n1_strong"):
    d=json.load(open(f"gpurun_out/{f}.json"))
    a=d["also"]
    print(f, round(d["value"]), d["ms_per_step"], d["roofline"]["frac"], d["parity"]["rel_l2"], "e2e", d["e2e"]["ms_per_step"], d["e2e"].get("pcie_floor_ms"))
    print(" 27pt", a["general_27pt_spmv"]["ms_per_step"], a["general_27pt_spmv"]["frac_of_hbm_peak"])
    for k in ("cg_27pt","cg_sym_7pt"):
        print(" ",k, a[k]["ms_per_iter_runs"], {x:round(y,1) for x,y in a[k]["phases"].items() if x!="how"}, a[k]["solve_rtol_1e-8"]["iterations"], round(a[k]["solve_rtol_1e-8"]["iters_per_s"]), a[k]["solve_rtol_1e-8"]["residual_true"])
    if "powerlaw_petsc" in a: print(" pl", a["powerlaw_petsc"]["ms_per_step"], a["powerlaw_petsc"]["local_kernel_frac_of_hbm_peak"])
    if "config1_createA" in a: print(" c1", a["config1_createA"]["us_per_step"], a["config1_createA"]["parity"]["bit_identical"])
    c=d.get("cpu_baseline")
    if c: print(" cpu", c["value"], c.get("ms_per_spmv"), c.get("ms_mean"), c.get("ms_min"))
    print(" clocks", d["clocks"], "launches", d["gpu_launches"])
PY
S0=$SECONDS
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r02_bench_ref_n1.json 2> gpurun_out/r02_bench_ref_n1.err; echo "ref rc=$? in $((SECONDS-S0)) s"
cut -c1-300 gpurun_out/r02_bench_ref_n1.json
