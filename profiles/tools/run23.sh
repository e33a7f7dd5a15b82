#!/bin/bash
# symmetric x'Ax epilogue: tests, then the bench (cg_sym_7pt carries the A/B)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_golden_gpu.py -x -q -m gpu -k "symv_dot or cg_symmetric_modes or cg_modes_agree or laplace_symmetric or random_symmetric" 2>&1 | tail -8 | cut -c1-300
S0=$SECONDS
timeout 1200 python bench.py > gpurun_out/r02_bench_n1_symdot.json 2> gpurun_out/r02_bench_n1_symdot.err; echo "bench rc=$? in $((SECONDS-S0)) s"
tail -3 gpurun_out/r02_bench_n1_symdot.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_bench_n1_symdot.json"))
a=d["also"]
print(round(d["value"]), d["ms_per_step"], d["roofline"]["frac"])
for k in ("cg_27pt","cg_sym_7pt"):
    print(k, a[k]["ms_per_iter_runs"], {x:round(y,1) for x,y in a[k]["phases"].items() if x!="how"}, a[k]["solve_rtol_1e-8"])
print(a["cg_sym_7pt"].get("fused_dot_ab"))
c=d["cpu_baseline"]; print("cpu", c["value"], c.get("ms_per_spmv"), c.get("ms_mean"), c.get("ms_min"))
PY
