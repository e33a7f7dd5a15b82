#!/bin/bash
# r02, 8 GPUs: distributed tests at 8 ranks, PCIe ceilings, weak + strong scaling records at N = 4, 8
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02_topo_n8.txt 2>&1
nproc > gpurun_out/r02_nproc.txt; free -g >> gpurun_out/r02_nproc.txt
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $3 "${@:4}"; }
tr 200 8 29508 profiles/tools/pcie_probe_n.py > gpurun_out/r02_pcie_n8_unbound.json 2> gpurun_out/r02_pcie.err
tr 200 8 29608 profiles/tools/pcie_probe_n.py --bind > gpurun_out/r02_pcie_n8_bound.json 2>> gpurun_out/r02_pcie.err
tr 200 4 29604 profiles/tools/pcie_probe_n.py --bind > gpurun_out/r02_pcie_n4_bound.json 2>> gpurun_out/r02_pcie.err
cat gpurun_out/r02_pcie_n8_unbound.json gpurun_out/r02_pcie_n8_bound.json | cut -c1-900
run() { tr 900 $1 $2 bench.py --gpus $1 --steps 20 --warmup 3 $3 > gpurun_out/$4.json 2> gpurun_out/$4.err; echo "$4 rc=$?"; tail -2 gpurun_out/$4.err | cut -c1-300; }
run 8 29701 "" r02_bench_n8_a
run 8 29702 "--strong --config5-rows 0" r02_bench_n8_strong_a
run 4 29703 "--strong --config5-rows 0" r02_bench_n4_strong_a
python - <<'PY'
import json
for f in ("r02_bench_n8_a","r02_bench_n8_strong_a","r02_bench_n4_strong_a"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
    except Exception as e:
        print(f, "no json", e); continue
    a=d["also"]
    print(f, round(d["value"]), d["ms_per_step"], d["parity"]["rel_l2"], "e2e", d["e2e"]["ms_per_step"], d["config"]["host"])
    print(" 27pt", a["general_27pt_spmv"]["ms_per_step"], a["general_27pt_spmv"]["parity"]["rel_l2"])
    for k in ("cg_27pt","cg_sym_7pt"):
        print(" ",k, a[k]["ms_per_iter_runs"], {x:round(y,1) for x,y in a[k]["phases"].items() if x!="how"}, a[k]["solve_rtol_1e-8"]["iterations"], round(a[k]["solve_rtol_1e-8"]["iters_per_s"]), a[k]["solve_rtol_1e-8"]["residual_true"])
    if "powerlaw_petsc" in a: print(" pl", a["powerlaw_petsc"]["ms_per_step"], a["powerlaw_petsc"]["local_kernel_frac_of_hbm_peak"], a["powerlaw_petsc"]["parity"]["max_rel_err"])
    if "config1_createA" in a: print(" c1", a["config1_createA"]["us_per_step"], a["config1_createA"]["parity"])
PY
timeout 400 python -m pytest tests/test_dist.py -x -q -m gpu > gpurun_out/r02_pytest_dist_n8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_dist_n8.log
tail -12 gpurun_out/r02_pytest_dist_n8.log | cut -c1-300
