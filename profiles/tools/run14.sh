#!/bin/bash
# r02: whole GPU suite + power-law A/B (tile order, evict_first) + short bench
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu_a.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu_a.log
tail -15 gpurun_out/r02_pytest_gpu_a.log
export KB_VARIANTS=8 KB_SHAPES=16384x3x16x2
for env in ""; do
  echo "== env: $env"
  env $env timeout 300 python profiles/tools/kbench.py --case powerlaw --n 3000000 --iters 20 2>&1 | grep -E "variant=|plan defaults"
done > gpurun_out/r02_powerlaw_split_ab.log 2>&1
cat gpurun_out/r02_powerlaw_split_ab.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_bench_n1_a.json 2> gpurun_out/r02_bench_n1_a.err; echo "bench rc=$?"
tail -3 gpurun_out/r02_bench_n1_a.err; python -c "
import json;d=json.load(open('gpurun_out/r02_bench_n1_a.json'))
print(json.dumps({k:d[k] for k in ('value','ms_per_step','parity','roofline')},indent=0)[:1500])
print(json.dumps(d['also'],indent=0)[:3000])"
